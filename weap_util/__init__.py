"""Drop-in for the reference's `weap_util` package (weap_util/weap_util/): same import path for the piece of it that sits
on the simulation step's downstream side, `weap_util.lidar.lidar_to_bitmap`, served by the CUDA rasteriser."""

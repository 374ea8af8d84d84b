"""`from weap_util.lidar import lidar_to_bitmap` (reference: weap_util/weap_util/lidar.py:114) -> red_gym_b200's CUDA rasteriser."""
from red_gym_b200.lidar_bitmap import LidarBitmap, lidar_to_bitmap   # noqa: F401

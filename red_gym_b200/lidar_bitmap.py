"""Batched scan -> bird's-eye bitmap on the GPU: the reference's `weap_util.lidar.lidar_to_bitmap`
(weap_util/weap_util/lidar.py:4-154) for N scans at once, byte-identical output.

    bm = LidarBitmap(num_beams=1080, draw_mode="FILL", bg_color="black")      # same keyword arguments as the reference
    images = bm(env.scans.view(-1, 1080))                                      # (n, 256, 256) uint8 CUDA tensor

The beam subsampling indices and the beam angle cos/sin tables are computed here with numpy exactly as the reference
computes them (np.linspace(..., dtype=int), np.cos); the per-scan work (polar -> pixel, polygon fill, lines, boxes)
runs in libf110_b200.so (csrc/f110_bitmap.cu).  No CPU path: without a CUDA device the constructor raises.
"""
import ctypes

import numpy as np

from . import _native as nat


def _tables(num_beams, target_beam_count, winding_dir, starting_angle, fov):
    direction = 1 if winding_dir == "CCW" else -1
    indices = np.linspace(0, num_beams - 1, target_beam_count, dtype=int)                     # lidar.py:66
    angles = starting_angle + direction * fov * np.linspace(0, 1, target_beam_count)          # lidar.py:70
    return (np.ascontiguousarray(indices, dtype=np.int32), np.ascontiguousarray(np.cos(angles)),
            np.ascontiguousarray(np.sin(angles)))


class LidarBitmap(object):
    """Keyword arguments and their defaults are those of `lidar_to_bitmap` (lidar.py:114-127); `num_beams` is len(scan),
    `scan_dtype` the dtype of the scan tensors ('fp64' like the reference env, or 'fp32' for the fast path)."""

    def __init__(self, num_beams=1080, winding_dir="CCW", starting_angle=-np.pi / 2, max_scan_radius=None, scaling_factor=10,
                 bg_color="white", draw_center=True, output_image_dims=(256, 256), target_beam_count=600, fov=2 * np.pi,
                 draw_mode="POLYGON", channels=1, scan_dtype="fp64", device=0):
        import torch
        self._torch = torch
        # the reference's argument checks (lidar.py:44-51,142), same exception types
        assert channels in [1, 3, 4], "channels must 1, 3, or 4"
        assert winding_dir in ["CW", "CCW"], "winding_dir must be either clockwise or counterclockwise"
        assert bg_color in ["black", "white"]
        assert draw_mode in ["RAYS", "POLYGON", "FILL"]
        assert len(output_image_dims) == 2
        assert all([x > 0 for x in output_image_dims]), "output_image_dims must be at least 1x1"
        assert 0 < target_beam_count < num_beams
        assert 0 < fov <= 2 * np.pi, "FOV must be between 0 and 2pi"
        if max_scan_radius is not None:
            scaling_factor = min(output_image_dims) / max_scan_radius
        elif scaling_factor is None:
            raise ValueError("Must provide either max_scan_radius or scaling_factor")
        if scan_dtype not in ("fp64", "fp32"):
            raise ValueError("scan_dtype must be 'fp64' or 'fp32'")
        self.num_beams, self.channels = int(num_beams), channels
        self.height, self.width = int(output_image_dims[0]), int(output_image_dims[1])
        self.scan_dtype = scan_dtype
        self.device = torch.device("cuda", int(device))
        self.lib = nat.load()
        idx, c, s = _tables(num_beams, target_beam_count, winding_dir, starting_angle, fov)
        cfg = nat.BitmapConfig(abi_version=nat.F110_ABI_VERSION, n_beams=self.num_beams, target_beam_count=int(target_beam_count),
                               height=self.height, width=self.width, draw_mode=nat.BITMAP_MODES[draw_mode],
                               draw_center=1 if draw_center else 0, bg_white=1 if bg_color == "white" else 0,
                               scan_is_f32=1 if scan_dtype == "fp32" else 0, scaling_factor=float(scaling_factor))
        h = ctypes.c_void_p()
        nat.check(self.lib.f110_bitmap_create(ctypes.byref(cfg), idx.ctypes.data_as(ctypes.c_void_p), c.ctypes.data_as(ctypes.c_void_p),
                                              s.ctypes.data_as(ctypes.c_void_p), int(device), ctypes.byref(h)))
        self._h = h

    def __call__(self, scans, out=None, packed=False):
        """scans: (n, num_beams) CUDA tensor of the handle's dtype -> (n, H, W) uint8 [or (n, H, W, channels)];
        packed=True: (n, H, ceil(W / 32)) int32 bit masks (bit set = drawn colour), 8x fewer bytes to write."""
        torch = self._torch
        want = torch.float64 if self.scan_dtype == "fp64" else torch.float32
        scans = torch.as_tensor(scans, device=self.device)
        if scans.dtype != want:
            raise TypeError("scan dtype %s does not match scan_dtype=%r" % (scans.dtype, self.scan_dtype))
        scans = scans.reshape(-1, self.num_beams).contiguous()
        n = scans.shape[0]
        if out is None:
            if packed:
                out = torch.empty((n, self.height, (self.width + 31) // 32), dtype=torch.int32, device=self.device)
            else:
                out = torch.empty((n, self.height, self.width), dtype=torch.uint8, device=self.device)
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        nat.check(self.lib.f110_bitmap_render(self._h, ctypes.c_void_p(scans.data_ptr()), n, ctypes.c_void_p(out.data_ptr()),
                                              1 if packed else 0, stream))
        if packed or self.channels == 1:
            return out
        if self.channels == 3:
            return out.unsqueeze(-1).expand(-1, -1, -1, 3)
        return torch.cat([out.unsqueeze(-1).expand(-1, -1, -1, 3), torch.full_like(out, 255).unsqueeze(-1)], dim=-1)

    def close(self):
        h = getattr(self, "_h", None)
        if h:
            self.lib.f110_bitmap_destroy(h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_cache = {}


def lidar_to_bitmap(scan, winding_dir="CCW", starting_angle=-np.pi / 2, max_scan_radius=None, scaling_factor=10,
                    bg_color="white", draw_center=True, output_image_dims=(256, 256), target_beam_count=600, fov=2 * np.pi,
                    draw_mode="POLYGON", channels=1):
    """Drop-in for weap_util.lidar.lidar_to_bitmap (lidar.py:114): one scan (list / numpy) in, numpy image out."""
    scan = np.array(scan)
    dt = "fp32" if scan.dtype == np.float32 else "fp64"
    key = (len(scan), winding_dir, float(starting_angle), max_scan_radius, scaling_factor, bg_color, bool(draw_center),
           tuple(output_image_dims), int(target_beam_count), float(fov), draw_mode, channels, dt)
    bm = _cache.get(key)
    if bm is None:
        bm = _cache[key] = LidarBitmap(len(scan), winding_dir, starting_angle, max_scan_radius, scaling_factor, bg_color,
                                       draw_center, output_image_dims, target_beam_count, fov, draw_mode, channels, dt)
    import torch
    t = torch.from_numpy(np.ascontiguousarray(scan if dt == "fp32" else scan.astype(np.float64)))[None].cuda()
    return bm(t)[0].cpu().numpy()

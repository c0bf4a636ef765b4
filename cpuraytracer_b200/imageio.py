"""Headless result sinks replacing the Direct2D preview (spheres-app.cpp:217-219): PPM (LDR) and PFM (HDR)."""
from __future__ import annotations

import numpy as np


def write_ppm(path: str, bgra8: np.ndarray) -> None:
    """bgra8: (H, W) uint32 0xAARRGGBB as XMStoreColor packs it (m_backbufferLdr, app.h:27)."""
    h, w = bgra8.shape
    rgb = np.empty((h, w, 3), np.uint8)
    rgb[..., 0] = (bgra8 >> 16) & 0xFF
    rgb[..., 1] = (bgra8 >> 8) & 0xFF
    rgb[..., 2] = bgra8 & 0xFF
    with open(path, "wb") as f:
        f.write(f"P6\n{w} {h}\n255\n".encode())
        f.write(rgb.tobytes())


def write_pfm(path: str, hdr_sum: np.ndarray, sample_count: int) -> None:
    """Linear mean radiance (sum / n) as a little-endian colour PFM (rows bottom-to-top per the format)."""
    h, w, _ = hdr_sum.shape
    mean = (hdr_sum[..., :3] / np.float32(max(sample_count, 1))).astype("<f4")
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1.0\n".encode())
        f.write(mean[::-1].tobytes())

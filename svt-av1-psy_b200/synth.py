"""Synthetic 4:2:0 luma clips for the ME benchmarks and tests: "panning noise".

A fixed-seed uint8 noise field, lightly low-passed, larger than the frame; frame t is the crop
at offset (vx*t, vy*t) plus small per-frame additive noise so SADs are non-degenerate and exact
ties are rare but present. 10-bit frames are the same field << 2 | 2 random LSBs."""
import numpy as np


def _lowpass(a):
    a = a.astype(np.float32)
    for axis in (0, 1):
        a = (np.roll(a, 1, axis) + 2 * a + np.roll(a, -1, axis)) * 0.25
    return a


class PanningNoise:
    def __init__(self, width, height, seed=7, vx=2, vy=1, max_frames=64, sigma=2.0, scale=4):
        self.w, self.h, self.vx, self.vy, self.sigma, self.seed = width, height, vx, vy, sigma, seed
        rng = np.random.default_rng(seed)
        fw = width + abs(vx) * max_frames + 8
        fh = height + abs(vy) * max_frames + 8
        # blocky base texture (structures at several scales) + fine noise
        coarse = rng.integers(0, 256, size=((fh + scale * 4 - 1) // (scale * 4), (fw + scale * 4 - 1) // (scale * 4)))
        coarse = np.kron(coarse, np.ones((scale * 4, scale * 4)))[:fh, :fw]
        fine = rng.integers(0, 256, size=(fh, fw))
        field = 0.6 * _lowpass(_lowpass(coarse)) + 0.4 * _lowpass(fine)
        self.field = np.clip(field, 0, 255).astype(np.uint8)
        self.max_frames = max_frames

    def frame(self, t):
        """8-bit luma of frame t, shape (h, w) uint8, C-contiguous."""
        tt = t % self.max_frames
        x0, y0 = abs(self.vx) * tt, abs(self.vy) * tt
        crop = self.field[y0:y0 + self.h, x0:x0 + self.w].astype(np.int16)
        rng = np.random.default_rng(self.seed * 1000003 + t)
        noise = np.rint(rng.normal(0.0, self.sigma, size=crop.shape)).astype(np.int16)
        return np.ascontiguousarray(np.clip(crop + noise, 0, 255).astype(np.uint8))

    def frame10(self, t):
        """10-bit luma of frame t as uint16 (8 MSBs equal frame(t))."""
        f8 = self.frame(t).astype(np.uint16)
        rng = np.random.default_rng(self.seed * 7919 + t)
        return np.ascontiguousarray((f8 << 2) | rng.integers(0, 4, size=f8.shape, dtype=np.uint16))

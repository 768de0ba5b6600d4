"""Parity gates shared by the model-level GPU tests (test_gpu_model.py, test_gpu_tp.py) and bench.py's parity leg.

Absolute gate (BASELINE.json north_star): max |dlogit| <= 1e-2.  With the frozen init the oracle's logits
have a standard deviation sigma of about 0.10, so the absolute gate alone is 10 % of sigma and a
regression that triples the error would pass.  Relative gates sit beside it.  They come from the CPU
calibration of the bf16 numerics contract (tools/calibrate_bf16_contract.py; table in BASELINE.md): the
error of ANY faithful bf16-operand pipeline against the fp32 oracle on this random-init model is
rms ~1.0 % / max ~5.4 % of sigma at 2 layers of Llama-3-8B width and grows like a random walk with depth
(x sqrt(L/2)): rms 4.0-4.2 % / max ~21 % of sigma at all 32 layers, i.e. max ~2e-2 ABSOLUTE.  The north-star
absolute bar is therefore attainable only up to ~6 layers of this width; for deeper models the gates below
scale with sqrt(depth / 2) and the absolute bar is replaced by the same relative maximum (stated, not hidden:
BASELINE.md "Parity at full depth").
"""
import math

import torch

LOGIT_TOL = 1e-2        # absolute, north_star (shallow models)
REL_MAX = 0.07          # max |dlogit| <= 0.07 sigma at depth <= 2
REL_RMS = 0.015         # rms dlogit  <= 0.015 sigma at depth <= 2
REL_MAX_DEEP = 0.09     # x sqrt(depth / 2): 0.36 sigma at 32 layers (calibration 0.21, GPU 0.28 over 17 M logits)


def gates_for(depth: int):
    """(absolute, max relative, rms relative) gates for a model of `depth` layers."""
    f = max(1.0, math.sqrt(depth / 2.0))
    rel_max = REL_MAX if f <= 1.0 else REL_MAX_DEEP * f
    return LOGIT_TOL if f <= 1.75 else None, rel_max, REL_RMS * f


class Gate:
    """Accumulates the worst absolute / relative error over the steps of one parity case."""

    def __init__(self, depth: int = 2):
        self.depth = depth
        self.max_abs = 0.0
        self.max_rel = 0.0
        self.rms_rel = 0.0
        self.sigma = 0.0

    def add(self, got: torch.Tensor, want: torch.Tensor):
        d = (got.float() - want.float())
        sigma = want.float().std().item()
        self.sigma = sigma
        self.max_abs = max(self.max_abs, d.abs().max().item())
        if sigma > 0:
            self.max_rel = max(self.max_rel, d.abs().max().item() / sigma)
            self.rms_rel = max(self.rms_rel, d.pow(2).mean().sqrt().item() / sigma)

    def summary(self) -> str:
        return (f"max|dlogit|={self.max_abs:.3e} ({self.max_rel * 100:.2f} % of sigma) rms={self.rms_rel * 100:.2f} % "
                f"of sigma, sigma={self.sigma:.3f}")

    def check(self, what: str = ""):
        abs_gate, rel_max, rel_rms = gates_for(self.depth)
        if abs_gate is not None:
            assert self.max_abs <= abs_gate, f"{what}: logits differ from the oracle by {self.max_abs} (> {abs_gate})"
        assert self.max_rel <= rel_max, f"{what}: max |dlogit| is {self.max_rel:.4f} sigma (> {rel_max:.3f} at depth {self.depth})"
        assert self.rms_rel <= rel_rms, f"{what}: rms dlogit is {self.rms_rel:.4f} sigma (> {rel_rms:.4f} at depth {self.depth})"

"""Parity gates shared by the model-level GPU tests (test_gpu_model.py, test_gpu_tp.py, bench.py's parity leg
reuses the numbers).

Absolute gate (BASELINE.json north_star): max |dlogit| <= 1e-2.  With the frozen init the oracle's logits
have a standard deviation sigma of about 0.10, so the absolute gate alone is 10 % of sigma and a
regression that triples the error would pass.  The relative gates sit beside it: they come from the CPU
calibration of the bf16 numerics contract in BASELINE.md ("Parity protocol as frozen in round 1": rms
0.6-1.0 % of sigma, maximum over B x V logits 3-5.5 % of sigma) with headroom for depth.
"""
import torch

LOGIT_TOL = 1e-2        # absolute, north_star
REL_MAX = 0.07          # max |dlogit| <= 0.07 sigma
REL_RMS = 0.015         # rms dlogit  <= 0.015 sigma


class Gate:
    """Accumulates the worst absolute / relative error over the steps of one parity case."""

    def __init__(self):
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
        assert self.max_abs <= LOGIT_TOL, f"{what}: logits differ from the oracle by {self.max_abs} (> {LOGIT_TOL})"
        assert self.max_rel <= REL_MAX, f"{what}: max |dlogit| is {self.max_rel:.4f} sigma (> {REL_MAX})"
        assert self.rms_rel <= REL_RMS, f"{what}: rms dlogit is {self.rms_rel:.4f} sigma (> {REL_RMS})"

"""Shared helpers for the parity tests."""
import glob
import os

import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# north_star tolerance for bf16 / fp8 outputs: max-abs <= 2e-2 and cosine >= 0.999 per layer
MAX_ABS_TOL = 2e-2
COS_TOL = 0.999


def golden_files(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN_DIR, f"{prefix}*.pt")))


def load_golden(path):
    return torch.load(path, map_location="cpu", weights_only=False)


def cosine(a: torch.Tensor, b: torch.Tensor) -> float:
    a = a.double().flatten()
    b = b.double().flatten()
    denom = a.norm() * b.norm()
    if denom == 0:
        return 1.0 if a.norm() == b.norm() else 0.0
    return float((a @ b) / denom)


def assert_close_path(actual: torch.Tensor, expected: torch.Tensor, what: str, max_abs=MAX_ABS_TOL, cos=COS_TOL):
    """The north_star tolerance: max-abs <= 2e-2 and cosine >= 0.999."""
    actual = actual.detach().cpu().float()
    expected = expected.detach().cpu().float()
    assert actual.shape == expected.shape, f"{what}: shape {tuple(actual.shape)} vs {tuple(expected.shape)}"
    assert torch.isfinite(actual).all(), f"{what}: non-finite values in the output"
    err = (actual - expected).abs().max().item() if actual.numel() else 0.0
    c = cosine(actual, expected) if actual.numel() else 1.0
    assert err <= max_abs and c >= cos, f"{what}: max-abs {err:.4e} (tol {max_abs}), cosine {c:.6f} (tol {cos})"
    return err, c

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


ROW_REL_TOL = 1e-2


def assert_close_path(actual: torch.Tensor, expected: torch.Tensor, what: str, max_abs=MAX_ABS_TOL, cos=COS_TOL,
                      row_rel=ROW_REL_TOL):
    """The north_star tolerance (max-abs <= 2e-2 and cosine >= 0.999), applied PER ROW (a row = one vector along the
    last dim: one (request, head) of an attention output, one token of a GEMM output), plus a per-row RELATIVE bound
    max|err| / max|row| <= row_rel.  A global cosine / absolute bound alone lets a long-context row — magnitude
    ~sqrt(1/len), well inside 2e-2 — lose a whole KV block unnoticed; the per-row relative bound does not."""
    actual = actual.detach().cpu().float()
    expected = expected.detach().cpu().float()
    assert actual.shape == expected.shape, f"{what}: shape {tuple(actual.shape)} vs {tuple(expected.shape)}"
    assert torch.isfinite(actual).all(), f"{what}: non-finite values in the output"
    if actual.numel() == 0:
        return 0.0, 1.0
    err = (actual - expected).abs().max().item()
    c = cosine(actual, expected)
    assert err <= max_abs and c >= cos, f"{what}: max-abs {err:.4e} (tol {max_abs}), cosine {c:.6f} (tol {cos})"
    a2 = actual.reshape(-1, actual.shape[-1]).double()
    e2 = expected.reshape(-1, expected.shape[-1]).double()
    row_max = e2.abs().amax(dim=1)
    floor = max(row_max.max().item() * 1e-3, 1e-30)  # rows that are (almost) zero are judged on the global scale
    rel = (a2 - e2).abs().amax(dim=1) / row_max.clamp_min(floor)
    worst = int(rel.argmax())
    assert rel[worst].item() <= row_rel, (f"{what}: row {worst} relative error {rel[worst].item():.4e} (tol {row_rel}; "
                                          f"row max-abs {row_max[worst].item():.4e})")
    na, ne = a2.norm(dim=1), e2.norm(dim=1)
    live = ne > floor
    if live.any():
        rc = ((a2 * e2).sum(dim=1)[live] / (na[live] * ne[live]).clamp_min(1e-300))
        w = int(rc.argmin())
        assert rc[w].item() >= cos, f"{what}: per-row cosine {rc[w].item():.6f} < {cos} (live row {w})"
    return err, c

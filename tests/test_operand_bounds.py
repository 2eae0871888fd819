"""CPU restatement of the operand arithmetic of the fused kernel (multimodal-fusion_b200/csrc/split.cu, knn_tc.cuh):
per-row power-of-two scale, FP16 hi / lo halves, and the error bounds `tc_eps` relies on -- checked numerically in
float64 on synthetic rows, no GPU.  (The GPU suite checks the kernel against the float64 oracle; this file checks that
the analytic bounds the candidate slack is built from hold for the split as specified.)"""
import numpy as np
import pytest

from hg_b200 import synth

U = 2.0 ** -11                      # FP16 unit roundoff


def op_scale_log2(sq_norm: np.ndarray) -> np.ndarray:
    """knn_tc.cuh op_scale_log2: |x|^2 in [2^(e-127), 2^(e-126)) -> s = 2^(6 - ceil((e - 126) / 2))"""
    e = (np.asarray(sq_norm, dtype=np.float32).view(np.int32) >> 23) & 0xFF
    return 6 - ((e - 125) >> 1)


def split_rows(x: np.ndarray):
    """hi = fp16(s x), lo = fp16(s x - hi) per row, as split.cu; returns float64 copies and the scales"""
    x = np.asarray(x, dtype=np.float32)
    norm = (x.astype(np.float64) ** 2).sum(1).astype(np.float32)
    s = np.ldexp(np.float32(1.0), op_scale_log2(norm)).astype(np.float32)
    xs = x * s[:, None]                                   # exact: a power of two
    hi = xs.astype(np.float16)
    res = xs - hi.astype(np.float32)                      # exact: the residue of a rounding
    lo = res.astype(np.float16)
    return xs.astype(np.float64), hi.astype(np.float64), lo.astype(np.float64), res.astype(np.float64), s


def test_row_scale_puts_every_norm_into_32_64():
    rng = np.random.default_rng(0)
    for d in (3, 64, 1536):
        x = rng.standard_normal((400, d)).astype(np.float32)
        x *= np.exp2(rng.integers(-60, 60, size=(400, 1))).astype(np.float32)        # norms over 36 orders of magnitude
        xs, hi, lo, res, s = split_rows(x)
        n = np.sqrt((xs ** 2).sum(1))
        assert (n >= 32.0 * (1 - 1e-6)).all() and (n < 64.0).all()
        assert np.abs(hi).max() < 65504 and np.isfinite(hi).all() and np.isfinite(lo).all()
    # a zero row and a denormal-norm row do not blow up
    z = np.zeros((2, 8), np.float32); z[1, 0] = 1e-30
    xs, hi, lo, res, s = split_rows(z)
    assert np.isfinite(s).all() and (hi[0] == 0).all()


@pytest.mark.parametrize("mode,d", [("gaussian", 1024), ("cosine", 200), ("gaussian", 1536)])
def test_hi_lo_halves_and_the_measured_residual(mode, d):
    x = synth.make_features(1500, d, 91, mode)
    xs, hi, lo, res, s = split_rows(x)
    # element-wise: |s x - hi| <= u |s x| (+ 2^-25 on the subnormal grid), hi + lo good to u^2 (+ 2^-25)
    assert (np.abs(res) <= U * np.abs(xs) + 2.0 ** -25).all()
    assert (np.abs(xs - hi - lo) <= U * U * np.abs(xs) + 2.0 ** -25).all()
    # rho as the split kernel measures it: at most u, and well below it because rounding errors do not line up
    rho_rows = np.sqrt((res ** 2).sum(1)) / np.sqrt((xs ** 2).sum(1))
    rho = rho_rows.max()
    assert rho <= U * (1 + 1e-6)
    assert rho < 0.6 * U, rho / U
    # two products (hi*hi + lo*hi: the B row rounded, the A row whole): error <= rho |a||b| by Cauchy-Schwarz
    rng = np.random.default_rng(5)
    ia, ib = rng.integers(0, len(x), 4000), rng.integers(0, len(x), 4000)
    na, nb = np.sqrt((xs[ia] ** 2).sum(1)), np.sqrt((xs[ib] ** 2).sum(1))
    exact = (xs[ia] * xs[ib]).sum(1)
    two = ((hi[ia] + lo[ia]) * hi[ib]).sum(1)
    err2 = np.abs(two - exact) / (na * nb)
    assert err2.max() <= rho + U * U + 2 * 2.0 ** -30 * np.sqrt(d)
    assert err2.max() <= U                                              # and inside the nominal worst case, of course
    # three products: only lo*lo and the rounding of lo are left out
    three = (hi[ia] * hi[ib] + hi[ia] * lo[ib] + lo[ia] * hi[ib]).sum(1)
    err3 = np.abs(three - exact) / (na * nb)
    assert err3.max() <= 3 * U * U + 2 * 2.0 ** -30 * np.sqrt(d)
    # one product (hi*hi, the optional coarse pre-pass): both rows rounded
    one = (hi[ia] * hi[ib]).sum(1)
    err1 = np.abs(one - exact) / (na * nb)
    assert err1.max() <= 2 * rho + U * U + 2 * 2.0 ** -30 * np.sqrt(d)


def test_dense_mode_halves_carry_19_bits():
    """dense-output operands: hi rounded to 8 significant bits first (a bf16 mantissa), lo = fp16 of the rest"""
    x = synth.make_features(600, 512, 92, "gaussian")
    norm = (x.astype(np.float64) ** 2).sum(1).astype(np.float32)
    s = np.ldexp(np.float32(1.0), op_scale_log2(norm)).astype(np.float32)
    xs = x * s[:, None]
    b = xs.view(np.uint32)
    hi8 = ((b + 0x7FFF + ((b >> 16) & 1)) & 0xFFFF0000).view(np.float32)          # round to nearest even at 8 bits
    hi = hi8.astype(np.float16).astype(np.float32)
    assert (hi == hi8).all()                                                      # an 8-bit mantissa fits an FP16 exactly here
    lo = (xs - hi).astype(np.float16).astype(np.float32)
    assert (np.abs(xs - hi) <= 2.0 ** -8 * np.abs(xs)).all()
    assert (np.abs(xs - hi - lo) <= 2.0 ** -19 * np.abs(xs) + 2.0 ** -25).all()

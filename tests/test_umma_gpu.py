"""tcgen05/TMEM GEMM (b200_op_w4_gemm_umma / b200_op_f16_gemm_umma) against the CPU oracle.
fp16 x fp16 products are exact in fp32 and the accumulator is fp32, so the tolerance is the
accumulation-order noise of an fp32 dot product (same bound as the CUDA-core GEMV)."""
import numpy as np
import pytest
import torch

from oracle.awq import dequantize_awq

pytestmark = pytest.mark.gpu

from tests.gpu_util import lib, dev, ptr, stream, check, repack, random_awq  # noqa: E402

TOL = 5e-4


def tile(wq, sc, zp, K, N, G=128, pair_rows=0):
    nbytes = lib().b200_op_w4_tile_bytes(K, N, G)
    assert nbytes == (N // 128) * (K // G) * (512 + (G // 64) * 4096)
    chunks = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    check(lib().b200_op_w4_tile(ptr(wq), ptr(sc), ptr(zp), K, N, G, pair_rows, ptr(chunks), stream()))
    return chunks


@pytest.mark.parametrize("T", [1, 16, 17, 64, 100, 200, 256, 300, 700])
@pytest.mark.parametrize("K,N", [(1536, 2048), (8960, 1536)])
def test_w4_umma_matches_oracle(T, K, N):
    rng = np.random.default_rng(T * 7 + K)
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    ch = tile(wq, s, z, K, N)
    x = rng.standard_normal((T, K)).astype(np.float16)
    bias = (rng.standard_normal(N) * 0.02).astype(np.float16)
    ref = x.astype(np.float64) @ dequantize_awq(qw, qz, sc, 128).astype(np.float64)
    xd, bd = dev(x), dev(bias)
    y = torch.full((T, N), 7.0, dtype=torch.float32, device="cuda")
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(y), N, ptr(bd), T, K, N, 128, 0, 1, 0,
                                     stream()))
    err = np.abs(y.cpu().numpy() - (ref + bias.astype(np.float64))).max()
    assert err < TOL, err
    # residual-accumulate epilogue
    y0 = rng.standard_normal((T, N)).astype(np.float32)
    y = dev(y0)
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(y), N, None, T, K, N, 128, 1, 1, 0,
                                     stream()))
    assert np.abs(y.cpu().numpy() - (y0 + ref)).max() < TOL


@pytest.mark.parametrize("splits", [2, 5, 12, 70])
def test_w4_umma_split_k_partials_sum_to_result(splits):
    rng = np.random.default_rng(splits)
    K, N, T = 8960, 1536, 64
    splits = lib().b200_op_w4_effective_splits(K, 128, splits)
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    ch = tile(wq, s, z, K, N)
    x = rng.standard_normal((T, K)).astype(np.float16)
    ref = x.astype(np.float64) @ dequantize_awq(qw, qz, sc, 128).astype(np.float64)
    parts = torch.zeros((splits, T, N), dtype=torch.float32, device="cuda")
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(dev(x)), K, ptr(parts), N, None, T, K, N, 128, 0,
                                     splits, T * N, stream()))
    got = parts.cpu().numpy().astype(np.float64).sum(0)
    assert np.abs(got - ref).max() < TOL


def test_umma_equals_gemv_kernel():
    """The two W4A16 kernels (CUDA-core GEMV, tcgen05 GEMM) agree to fp32 accumulation noise."""
    rng = np.random.default_rng(9)
    K, N, T = 1536, 2048, 8
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    ch = tile(wq, s, z, K, N)
    xd = dev(rng.standard_normal((T, K)).astype(np.float16))
    y1 = torch.empty((T, N), dtype=torch.float32, device="cuda")
    y2 = torch.empty_like(y1)
    check(lib().b200_op_w4_gemv(ptr(wq), ptr(s), ptr(z), ptr(xd), K, ptr(y1), N, None, T, K, N, 128, 0, stream()))
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(y2), N, None, T, K, N, 128, 0, 1, 0,
                                     stream()))
    assert (y1 - y2).abs().max().item() < 1e-4


@pytest.mark.parametrize("T", [1, 40, 64, 256, 260])
def test_f16_umma_matches_oracle(T):
    rng = np.random.default_rng(T)
    K, N = 1536, 8192
    w = (rng.standard_normal((N, K)) * 0.02).astype(np.float16)
    x = rng.standard_normal((T, K)).astype(np.float16)
    y = torch.empty((T, N), dtype=torch.float32, device="cuda")
    check(lib().b200_op_f16_gemm_umma(ptr(dev(w)), ptr(dev(x)), K, ptr(y), N, T, K, N, 0, stream()))
    ref = x.astype(np.float64) @ w.astype(np.float64).T
    assert np.abs(y.cpu().numpy() - ref).max() < TOL


@pytest.mark.parametrize("T", [1, 17, 64, 130])
def test_lm_head_argmax_equals_argmax_of_logits(T):
    """Fused greedy lm_head: bit-identical choice to argmax over the logits the store epilogue writes
    (first maximum wins), including exact ties planted in different 128-row tiles."""
    rng = np.random.default_rng(100 + T)
    K, N = 1536, 8192
    w = (rng.standard_normal((N, K)) * 0.02).astype(np.float16)
    w[4000] = w[77]                                  # equal logits in tiles 0 and 31 -> the lower index must win
    w[77 + 128] = w[77]
    x = rng.standard_normal((T, K)).astype(np.float16)
    x[0] = (w[77].astype(np.float32) * 40).astype(np.float16)       # row 0 points at the tied rows
    wd, xd = dev(w), dev(x)
    y = torch.empty((T, N), dtype=torch.float32, device="cuda")
    check(lib().b200_op_f16_gemm_umma(ptr(wd), ptr(xd), K, ptr(y), N, T, K, N, 0, stream()))
    scratch = torch.empty((N // 128, T, 2), dtype=torch.float32, device="cuda")
    out = torch.full((T,), -1, dtype=torch.int32, device="cuda")
    check(lib().b200_op_lm_head_argmax(ptr(wd), ptr(xd), K, T, K, N, ptr(scratch), ptr(out), stream()))
    logits = y.cpu().numpy()
    assert np.array_equal(out.cpu().numpy(), logits.argmax(-1).astype(np.int32))     # numpy: first maximum
    assert out[0].item() == 77


def test_w4_umma_group_size_64():
    rng = np.random.default_rng(64)
    K, N, T, G = 1536, 256, 33, 64
    qw, qz, sc = random_awq(rng, K, N, G=G)
    wq, s, z = repack(qw, qz, sc, G=G)
    ch = tile(wq, s, z, K, N, G=G)
    x = rng.standard_normal((T, K)).astype(np.float16)
    y = torch.empty((T, N), dtype=torch.float32, device="cuda")
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(dev(x)), K, ptr(y), N, None, T, K, N, G, 0, 1, 0, stream()))
    ref = x.astype(np.float64) @ dequantize_awq(qw, qz, sc, G).astype(np.float64)
    assert np.abs(y.cpu().numpy() - ref).max() < TOL


def test_w4_umma_is_batch_invariant():
    """A token's output bits do not depend on which other tokens share the launch (tile height,
    position in the tile) -- the property continuous batching relies on."""
    rng = np.random.default_rng(11)
    K, N = 1536, 2048
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    ch = tile(wq, s, z, K, N)
    x = rng.standard_normal((300, K)).astype(np.float16)
    xd = dev(x)
    outs = {}
    for T in (1, 5, 64, 300):
        y = torch.empty((T, N), dtype=torch.float32, device="cuda")
        check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(y), N, None, T, K, N, 128, 0, 1, 0, stream()))
        outs[T] = y.cpu().numpy()
    for T in (5, 64, 300):
        assert np.array_equal(outs[T][0], outs[1][0])
    assert np.array_equal(outs[300][:64], outs[64])


@pytest.mark.parametrize("T", [1, 6, 64, 100, 300])
def test_w4_umma_fused_swiglu(T):
    """gate_up GEMM with the SwiGLU epilogue == silu(x Wg) * (x Wu), one fp16 rounding."""
    rng = np.random.default_rng(T + 5)
    K, I = 1536, 1280
    qw, qz, sc = random_awq(rng, K, 2 * I)                 # columns [gate | up]
    wq, s, z = repack(qw, qz, sc)
    ch = tile(wq, s, z, K, 2 * I, pair_rows=I)
    x = rng.standard_normal((T, K)).astype(np.float16)
    act = torch.zeros((T, I), dtype=torch.float16, device="cuda")
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(dev(x)), K, ptr(act), I, None, T, K, 2 * I, 128, 2, 1, 0, stream()))
    gu = x.astype(np.float64) @ dequantize_awq(qw, qz, sc, 128).astype(np.float64)
    g, u = gu[:, :I], gu[:, I:]
    ref = g / (1 + np.exp(-g)) * u
    got = act.cpu().numpy().astype(np.float64)
    assert (np.abs(got - ref) <= 1.5e-3 * np.abs(ref) + 1e-4).all()


def test_w4_umma_two_row_tiles_per_cta_bitwise():
    """Large prefill passes run 256 weight rows per CTA (two accumulators sharing the staged activations).
    Its bits must equal the 128-row kernel's: the first 128 tokens of a 1280-token call (256-row path: N/256 x
    10 token tiles >= 296 CTAs) against a 128-token call of the same rows (128-row path) -- store + bias,
    accumulate, and the fused SwiGLU epilogue -- plus a check of the whole result against the oracle."""
    rng = np.random.default_rng(77)
    K, N, T = 1536, 8192, 1280
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    ch = tile(wq, s, z, K, N)
    chp = tile(wq, s, z, K, N, pair_rows=N // 2)
    x = rng.standard_normal((T, K)).astype(np.float16)
    bias = (rng.standard_normal(N) * 0.1).astype(np.float16)
    xd, bd = dev(x), dev(bias)
    big = torch.empty((T, N), dtype=torch.float32, device="cuda")
    small = torch.empty((128, N), dtype=torch.float32, device="cuda")
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(big), N, ptr(bd), T, K, N, 128, 0, 1, 0, stream()))
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(small), N, ptr(bd), 128, K, N, 128, 0, 1, 0, stream()))
    assert torch.equal(big[:128].view(torch.int32), small.view(torch.int32))
    ref = x[:256].astype(np.float64) @ dequantize_awq(qw, qz, sc, 128).astype(np.float64) + bias.astype(np.float64)
    assert np.abs(big[:256].cpu().numpy() - ref).max() < TOL
    # accumulate
    y0 = rng.standard_normal((T, N)).astype(np.float32)
    yb, ys = dev(y0), dev(y0[:128].copy())
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(yb), N, None, T, K, N, 128, 1, 1, 0, stream()))
    check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(ys), N, None, 128, K, N, 128, 1, 1, 0, stream()))
    assert torch.equal(yb[:128].view(torch.int32), ys.view(torch.int32))
    # fused SwiGLU (rows paired per tile)
    ab = torch.zeros((T, N // 2), dtype=torch.float16, device="cuda")
    as_ = torch.zeros((128, N // 2), dtype=torch.float16, device="cuda")
    check(lib().b200_op_w4_gemm_umma(ptr(chp), ptr(xd), K, ptr(ab), N // 2, None, T, K, N, 128, 2, 1, 0, stream()))
    check(lib().b200_op_w4_gemm_umma(ptr(chp), ptr(xd), K, ptr(as_), N // 2, None, 128, K, N, 128, 2, 1, 0, stream()))
    assert torch.equal(ab[:128].view(torch.int16), as_.view(torch.int16))
    assert ab[1000:].float().abs().sum().item() > 0

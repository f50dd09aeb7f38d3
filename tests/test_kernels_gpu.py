"""Parity of every kernel on the path against the CPU oracle, through the C ABI (b200_op_*).
Integer / index work is bit-exact; floating point within the stated tolerance."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle.awq import unpack_awq, dequantize_awq
from oracle.qwen2 import Qwen2Dims, rope_table
from oracle import sampler as osampler

pytestmark = pytest.mark.gpu

from tests.gpu_util import lib, dev, ptr, stream, check, repack, random_awq, swizzle_kv, unswizzle_kv  # noqa: E402


@pytest.mark.parametrize("K,N", [(1536, 2048), (8960, 1536), (256, 64)])
def test_awq_unpack_bit_exact(K, N):
    rng = np.random.default_rng(K + N)
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    # integer planes: kernel layout re-read on the host must equal the checkpoint's nibbles
    q_ref = unpack_awq(qw)                                  # [K, N]
    got = unpack_awq(wq.cpu().numpy().reshape(N, K // 8))   # same nibble order, along K -> [N, K]
    assert np.array_equal(got, q_ref.T)
    assert np.array_equal(z.cpu().numpy().astype(np.int32) - 1024, unpack_awq(qz).T.astype(np.int32))
    assert np.array_equal(s.cpu().numpy(), sc.T)
    # dequantised fp16 weights bit-exact against the reference formula fp16((q - z) * s)
    w = torch.empty((N, K), dtype=torch.float16, device="cuda")
    check(lib().b200_op_awq_dequant(ptr(wq), ptr(s), ptr(z), K, N, 128, ptr(w), stream()))
    ref = dequantize_awq(qw, qz, sc, 128)                   # [K, N] fp16
    assert np.array_equal(w.cpu().numpy().view(np.uint16), ref.T.copy().view(np.uint16))


@pytest.mark.parametrize("B", [1, 2, 3, 8, 13])
@pytest.mark.parametrize("K,N", [(1536, 2048), (8960, 1536)])
def test_w4_gemv_matches_oracle(B, K, N):
    rng = np.random.default_rng(B * 1000 + K)
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    x = (rng.standard_normal((B, K)).astype(np.float32)).astype(np.float16)
    bias = (rng.standard_normal(N) * 0.02).astype(np.float16)
    y0 = rng.standard_normal((B, N)).astype(np.float32)
    w_ref = dequantize_awq(qw, qz, sc, 128).astype(np.float64)
    ref = x.astype(np.float64) @ w_ref + bias.astype(np.float64)
    xd, bd = dev(x), dev(bias)
    y = torch.empty((B, N), dtype=torch.float32, device="cuda")
    check(lib().b200_op_w4_gemv(ptr(wq), ptr(s), ptr(z), ptr(xd), K, ptr(y), N, ptr(bd), B, K, N, 128, 0, stream()))
    assert np.abs(y.cpu().numpy() - ref).max() < 2e-4          # fp32 accumulation of exact products
    y = dev(y0)
    check(lib().b200_op_w4_gemv(ptr(wq), ptr(s), ptr(z), ptr(xd), K, ptr(y), N, None, B, K, N, 128, 1, stream()))
    assert np.abs(y.cpu().numpy() - (y0 + ref - bias.astype(np.float64))).max() < 2e-4


@pytest.mark.parametrize("B", [1, 4, 9])
def test_f16_gemv_matches_oracle(B):
    rng = np.random.default_rng(B)
    K, N = 1536, 4096
    w = (rng.standard_normal((N, K)) * 0.02).astype(np.float16)
    x = rng.standard_normal((B, K)).astype(np.float16)
    wd, xd = dev(w), dev(x)
    y = torch.empty((B, N), dtype=torch.float32, device="cuda")
    check(lib().b200_op_f16_gemv(ptr(wd), ptr(xd), K, ptr(y), N, B, K, N, 0, stream()))
    ref = x.astype(np.float64) @ w.astype(np.float64).T
    assert np.abs(y.cpu().numpy() - ref).max() < 2e-4


def test_rmsnorm_with_partials_and_row_map():
    rng = np.random.default_rng(0)
    T, H, P = 7, 1536, 3
    h = rng.standard_normal((T, H)).astype(np.float32)
    parts = rng.standard_normal((P, T, H)).astype(np.float32)
    g = (1 + 0.1 * rng.standard_normal(H)).astype(np.float16)
    rows = np.array([6, 0, 3], dtype=np.int32)
    hd, pd, gd, rd = dev(h), dev(parts), dev(g), dev(rows)
    xn = torch.empty((3, H), dtype=torch.float16, device="cuda")
    check(lib().b200_op_rmsnorm(ptr(hd), ptr(pd), P, T * H, ptr(gd), 1e-6, ptr(rd), 3, H, ptr(xn), stream()))
    v = h.astype(np.float64) + parts.astype(np.float64).sum(0)
    ref = v[rows] / np.sqrt((v[rows] ** 2).mean(-1, keepdims=True) + 1e-6) * g.astype(np.float64)
    got = xn.cpu().numpy().astype(np.float64)
    assert np.abs(got - ref).max() < 4e-3                       # one fp16 rounding of O(1..4) values
    # residual stream updated in place only for the rows that were visited
    h_after = hd.cpu().numpy()
    assert np.abs(h_after[rows] - v[rows]).max() < 1e-5
    assert np.array_equal(h_after[[1, 2, 4, 5]], h[[1, 2, 4, 5]])


@pytest.mark.parametrize("H", [1536, 256, 2048])
def test_rmsnorm_prefill_variant_is_bit_identical(H):
    """Large passes without partials use a warp-per-row kernel; it must produce the bits of the cluster kernel."""
    rng = np.random.default_rng(5)
    T = 1500
    h = (rng.standard_normal((T, H)) * 3).astype(np.float32)
    g = (1 + 0.1 * rng.standard_normal(H)).astype(np.float16)
    hd, gd = dev(h), dev(g)
    big = torch.empty((T, H), dtype=torch.float16, device="cuda")
    check(lib().b200_op_rmsnorm(ptr(hd), None, 0, 0, ptr(gd), 1e-6, None, T, H, ptr(big), stream()))        # >= 1024 rows
    small = torch.empty((T, H), dtype=torch.float16, device="cuda")
    for r0 in (0, 750):                                                                                        # < 1024 rows each
        check(lib().b200_op_rmsnorm(C.c_void_p(hd.data_ptr() + r0 * H * 4), None, 0, 0, ptr(gd), 1e-6, None, 750, H,
                                    C.c_void_p(small.data_ptr() + r0 * H * 2), stream()))
    assert torch.equal(big.view(torch.int16), small.view(torch.int16))
    v = h.astype(np.float64)
    ref = v / np.sqrt((v ** 2).mean(-1, keepdims=True) + 1e-6) * g.astype(np.float64)
    assert np.abs(big.cpu().numpy().astype(np.float64) - ref).max() < 2e-2


def test_embed_swiglu():
    rng = np.random.default_rng(1)
    V, H, T, I = 1000, 1536, 5, 8960
    tab = (rng.standard_normal((V, H)) * 0.02).astype(np.float16)
    ids = rng.integers(0, V, T).astype(np.int32)
    h = torch.empty((T, H), dtype=torch.float32, device="cuda")
    check(lib().b200_op_embed(ptr(dev(tab)), ptr(dev(ids)), T, H, ptr(h), stream()))
    assert np.array_equal(h.cpu().numpy(), tab[ids].astype(np.float32))
    gu = rng.standard_normal((T, 2 * I)).astype(np.float32) * 2
    act = torch.empty((T, I), dtype=torch.float16, device="cuda")
    check(lib().b200_op_swiglu(ptr(dev(gu)), T, I, ptr(act), stream()))
    g, u = gu[:, :I].astype(np.float64), gu[:, I:].astype(np.float64)
    ref = g / (1 + np.exp(-g)) * u
    got = act.cpu().numpy().astype(np.float64)
    assert (np.abs(got - ref) <= 1e-3 * np.abs(ref) + 1e-5).all()     # one fp16 rounding (rel 2^-11)


def _paged_setup(rng, n_seqs, ctx_lens, n_kv, n_blocks):
    """Random K/V written into randomly assigned blocks; returns caches, block tables, dense copies."""
    max_blocks = (max(ctx_lens) + 15) // 16 + 1
    perm = rng.permutation(n_blocks)
    bt = np.zeros((n_seqs, max_blocks), dtype=np.int32)
    kc = (rng.standard_normal((n_blocks, n_kv, 16, 128))).astype(np.float16)   # garbage everywhere
    vc = (rng.standard_normal((n_blocks, n_kv, 16, 128))).astype(np.float16)
    dense_k, dense_v, used = [], [], 0
    for s, L in enumerate(ctx_lens):
        nb = (L + 15) // 16
        bt[s, :nb] = perm[used:used + nb]
        used += nb
        k = np.concatenate([kc[bt[s, b]] for b in range(nb)], axis=1)[:, :L]   # [n_kv, L, 128]
        v = np.concatenate([vc[bt[s, b]] for b in range(nb)], axis=1)[:, :L]
        dense_k.append(k)
        dense_v.append(v)
    return kc, vc, bt, dense_k, dense_v


def _attn_ref(q, dk, dv, n_heads, n_kv):
    grp = n_heads // n_kv
    out = np.zeros((len(dk), n_heads, 128))
    for s in range(len(dk)):
        for h in range(n_heads):
            k, v = dk[s][h // grp].astype(np.float64), dv[s][h // grp].astype(np.float64)
            sc = k @ q[s, h].astype(np.float64) / np.sqrt(128.0)
            p = np.exp(sc - sc.max())
            out[s, h] = (p / p.sum()) @ v
    return out


@pytest.mark.parametrize("n_splits", [1, 3, 8])
@pytest.mark.parametrize("n_heads,n_kv", [(12, 2), (28, 4), (4, 4)])
def test_paged_attention_matches_oracle(n_heads, n_kv, n_splits):
    rng = np.random.default_rng(n_heads)
    ctx_lens = [1, 15, 16, 17, 64, 65, 200, 333, 0, 640]
    n_seqs = len(ctx_lens)
    kc, vc, bt, dk, dv = _paged_setup(rng, n_seqs, [max(c, 1) for c in ctx_lens], n_kv, n_blocks=128)
    q = rng.standard_normal((n_seqs, n_heads, 128)).astype(np.float16)
    out = torch.full((n_seqs, n_heads * 128), 7.0, dtype=torch.float16, device="cuda")
    scratch = torch.empty(lib().b200_op_attn_scratch_bytes(n_seqs, n_heads, n_kv, n_splits) // 4, dtype=torch.float32,
                          device="cuda")
    counters = torch.zeros(n_seqs * n_kv, dtype=torch.int32, device="cuda")
    for rep in range(2):                # twice: the arrival counters must clean up after themselves
        check(lib().b200_op_paged_attn(ptr(dev(q)), ptr(dev(swizzle_kv(kc))), ptr(dev(swizzle_kv(vc))), ptr(dev(bt)), bt.shape[1],
                                       ptr(dev(np.arange(n_seqs, dtype=np.int32))),
                                       ptr(dev(np.array(ctx_lens, dtype=np.int32))), n_seqs, n_heads, n_kv, n_splits,
                                       ptr(scratch), ptr(counters), ptr(out), stream()))
        assert not counters.any().item()
    got = out.cpu().numpy().astype(np.float64).reshape(n_seqs, n_heads, 128)
    ref = _attn_ref(q, dk, dv, n_heads, n_kv)
    for s in range(n_seqs):
        if ctx_lens[s] == 0:
            assert not got[s].any()              # empty context: zeros
        else:
            assert np.abs(got[s] - ref[s]).max() < 4e-3, s          # P is rounded to fp16 for the tensor-core PV


@pytest.mark.parametrize("n_splits,n_parts", [(1, 1), (8, 4)])
def test_fused_decode_attention_matches_rope_plus_attention(n_splits, n_parts):
    """The fused decode kernel == b200_op_rope_kv followed by b200_op_paged_attn: same q rotation, same
    cache contents afterwards, same attention output."""
    rng = np.random.default_rng(17 + n_splits)
    n_heads, n_kv, n_blocks = 12, 2, 96
    ctx_lens = [1, 16, 17, 64, 65, 129, 300, 512]           # context INCLUDING the new token
    T = len(ctx_lens)
    kc, vc, bt, _, _ = _paged_setup(rng, T, ctx_lens, n_kv, n_blocks)
    pos = np.array([c - 1 for c in ctx_lens], dtype=np.int32)
    slot = np.array([bt[i, p // 16] * 16 + p % 16 for i, p in enumerate(pos)], dtype=np.int32)
    cos, sin = rope_table(Qwen2Dims(), 4096)
    width = (n_heads + 2 * n_kv) * 128
    parts = rng.standard_normal((n_parts, T, width)).astype(np.float32)
    bias = (rng.standard_normal(width) * 0.05).astype(np.float16)
    btrow, ctx = np.arange(T, dtype=np.int32), np.array(ctx_lens, dtype=np.int32)
    cosd, sind, btd, posd, slotd, partsd, biasd = dev(cos.numpy()), dev(sin.numpy()), dev(bt), dev(pos), dev(slot), dev(parts), dev(bias)
    btrowd, ctxd = dev(btrow), dev(ctx)
    # reference path: rope kernel then plain attention, on its own copy of the cache
    kc1, vc1 = dev(swizzle_kv(kc)).clone(), dev(swizzle_kv(vc)).clone()
    qo = torch.empty((T, n_heads * 128), dtype=torch.float16, device="cuda")
    out1 = torch.empty_like(qo)
    check(lib().b200_op_rope_kv(ptr(partsd), n_parts, T * width, ptr(biasd), ptr(posd), ptr(slotd), ptr(cosd), ptr(sind), T,
                                n_heads, n_kv, ptr(qo), ptr(kc1), ptr(vc1), stream()))
    cosd, sind, btd, posd, slotd, partsd, biasd = dev(cos.numpy()), dev(sin.numpy()), dev(bt), dev(pos), dev(slot), dev(parts), dev(bias)
    btrowd, ctxd = dev(btrow), dev(ctx)
    check(lib().b200_op_paged_attn(ptr(qo), ptr(kc1), ptr(vc1), ptr(btd), bt.shape[1], ptr(btrowd), ptr(ctxd), T, n_heads,
                                   n_kv, 1, None, None, ptr(out1), stream()))
    # fused path
    kc2, vc2 = dev(swizzle_kv(kc)).clone(), dev(swizzle_kv(vc)).clone()
    out2 = torch.empty_like(qo)
    scratch = torch.empty(lib().b200_op_attn_scratch_bytes(T, n_heads, n_kv, n_splits) // 4, dtype=torch.float32, device="cuda")
    counters = torch.zeros(T * n_kv, dtype=torch.int32, device="cuda")
    cosd, sind, btd, posd, slotd, partsd, biasd = dev(cos.numpy()), dev(sin.numpy()), dev(bt), dev(pos), dev(slot), dev(parts), dev(bias)
    btrowd, ctxd = dev(btrow), dev(ctx)
    check(lib().b200_op_attn_decode_fused(ptr(partsd), n_parts, T * width, ptr(biasd), ptr(posd), ptr(slotd), ptr(cosd),
                                          ptr(sind), ptr(kc2), ptr(vc2), ptr(btd), bt.shape[1], ptr(btrowd), ptr(ctxd), T,
                                          n_heads, n_kv, n_splits, ptr(scratch), ptr(counters), ptr(out2), stream()))
    assert torch.equal(kc1, kc2) and torch.equal(vc1, vc2)              # cache writes are bit-identical
    assert not counters.any().item()
    d = (out1.float() - out2.float()).abs().max().item()
    assert d < (1e-6 if n_splits == 1 else 2e-3), d                      # merge order differs only when split


def test_fused_decode_attention_two_ctas_per_sm_variant_bitwise():
    """With more CTAs than SMs the decode attention runs three tile-owning warps (two CTAs per SM); the chunks
    are still dealt to four virtual warps, so its bits equal the four-warp kernel's: 100 rows in one call
    (200 CTAs -> three-warp kernel) against the same rows in two calls of 50 (four-warp kernel)."""
    rng = np.random.default_rng(31)
    n_heads, n_kv, n_blocks = 12, 2, 2700
    ctx_lens = [int(x) for x in rng.integers(1, 400, size=100)]
    ctx_lens[:6] = [1, 64, 65, 192, 193, 257]
    T = len(ctx_lens)
    kc, vc, bt, _, _ = _paged_setup(rng, T, ctx_lens, n_kv, n_blocks)
    pos = np.array([c - 1 for c in ctx_lens], dtype=np.int32)
    slot = np.array([bt[i, p // 16] * 16 + p % 16 for i, p in enumerate(pos)], dtype=np.int32)
    cos, sin = rope_table(Qwen2Dims(), 4096)
    width = (n_heads + 2 * n_kv) * 128
    n_parts = 4
    parts = rng.standard_normal((n_parts, T, width)).astype(np.float32)
    bias = (rng.standard_normal(width) * 0.05).astype(np.float16)
    cosd, sind, btd, biasd = dev(cos.numpy()), dev(sin.numpy()), dev(bt), dev(bias)
    ctx = np.array(ctx_lens, dtype=np.int32)

    def run(rows):
        n = len(rows)
        kcd, vcd = dev(swizzle_kv(kc)).clone(), dev(swizzle_kv(vc)).clone()
        out = torch.empty((n, n_heads * 128), dtype=torch.float16, device="cuda")
        pd, sd, cd = dev(pos[rows]), dev(slot[rows]), dev(ctx[rows])
        brd = dev(np.asarray(rows, dtype=np.int32))
        prt = dev(np.ascontiguousarray(parts[:, rows]))
        check(lib().b200_op_attn_decode_fused(ptr(prt), n_parts, n * width, ptr(biasd), ptr(pd), ptr(sd), ptr(cosd), ptr(sind),
                                              ptr(kcd), ptr(vcd), ptr(btd), bt.shape[1], ptr(brd), ptr(cd), n, n_heads, n_kv,
                                              1, None, None, ptr(out), stream()))
        return out, kcd, vcd

    rows = list(range(T))
    big, kb, vb = run(rows)
    lo, k1, v1 = run(rows[:50])
    hi, k2, v2 = run(rows[50:])
    assert torch.equal(big[:50].view(torch.int16), lo.view(torch.int16))
    assert torch.equal(big[50:].view(torch.int16), hi.view(torch.int16))


def test_rope_kv_write_matches_oracle():
    rng = np.random.default_rng(3)
    n_heads, n_kv, T, n_blocks = 12, 2, 9, 8
    dims = Qwen2Dims()
    cos, sin = rope_table(dims, 4096)
    qkv = rng.standard_normal((T, (n_heads + 2 * n_kv) * 128)).astype(np.float32)
    pos = np.array([0, 1, 5, 17, 100, 2047, 3, 3, 4095], dtype=np.int32)
    slot = np.array([0, 1, 37, 16, 127, 64, -1, 90, 5], dtype=np.int32)
    kc = torch.zeros((n_blocks, n_kv, 16, 128), dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    qo = torch.empty((T, n_heads * 128), dtype=torch.float16, device="cuda")
    # the row arrives as 3 split-K partials + bias, summed inside the kernel
    bias = (rng.standard_normal(qkv.shape[1]) * 0.02).astype(np.float16)
    parts = rng.standard_normal((3,) + qkv.shape).astype(np.float32)
    parts[2] = qkv - bias.astype(np.float32) - parts[0] - parts[1]
    qkv = (parts[0] + parts[1]) + parts[2] + bias.astype(np.float32)        # the kernel's summation order
    check(lib().b200_op_rope_kv(ptr(dev(parts)), 3, qkv.size, ptr(dev(bias)), ptr(dev(pos)), ptr(dev(slot)),
                                ptr(dev(cos.numpy())), ptr(dev(sin.numpy())), T, n_heads, n_kv, ptr(qo), ptr(kc), ptr(vc),
                                stream()))

    def rot(x, p):                      # x [..., 128]
        c, s = cos[p].numpy().astype(np.float64), sin[p].numpy().astype(np.float64)
        x1, x2 = x[..., :64].astype(np.float64), x[..., 64:].astype(np.float64)
        return np.concatenate([x1 * c - x2 * s, x2 * c + x1 * s], -1)

    kcn, vcn = unswizzle_kv(kc.cpu().numpy()), unswizzle_kv(vc.cpu().numpy())      # physical rows -> logical
    qon = qo.cpu().numpy().reshape(T, n_heads, 128)
    for t in range(T):
        row = qkv[t].reshape(n_heads + 2 * n_kv, 128)
        assert np.abs(qon[t].astype(np.float64) - rot(row[:n_heads], pos[t])).max() < 4e-3
        if slot[t] < 0:
            continue
        b, o = slot[t] // 16, slot[t] % 16
        assert np.abs(kcn[b, :, o].astype(np.float64) - rot(row[n_heads:n_heads + n_kv], pos[t])).max() < 4e-3
        assert np.array_equal(vcn[b, :, o], row[n_heads + n_kv:].astype(np.float16))
    # nothing else was written
    written = np.zeros((n_blocks, 16), bool)
    for s_ in slot[slot >= 0]:
        written[s_ // 16, s_ % 16] = True
    assert not kcn.transpose(0, 2, 1, 3)[~written].any() and not vcn.transpose(0, 2, 1, 3)[~written].any()


def test_sampler_greedy_and_top_p():
    rng = np.random.default_rng(4)
    B, V = 12, 151936
    logits = (rng.standard_normal((B, V)) * 2.5).astype(np.float32)
    logits[3, 777] = logits[3].max()          # exact tie: lowest index must win
    temp = np.array([0, 0, 0, 0, 0.7, 0.7, 1.0, 1.0, 0.3, 1.5, 0.7, 0.7], dtype=np.float32)
    topp = np.array([1, .9, .5, 1, .9, .5, .9, 1.0, .95, .8, .1, .99], dtype=np.float32)
    seed = rng.integers(0, 2**32, B, dtype=np.uint32)
    step = rng.integers(0, 300, B).astype(np.int32)
    out = torch.empty(B, dtype=torch.int32, device="cuda")
    check(lib().b200_op_sample(ptr(dev(logits)), B, V, ptr(dev(temp)), ptr(dev(topp)), ptr(dev(seed.view(np.int32))),
                               ptr(dev(step)), ptr(out), stream()))
    got = out.cpu().numpy()
    n_same = 0
    for r in range(B):
        ref = osampler.sample(logits[r], float(temp[r]), float(topp[r]), int(seed[r]), int(step[r]))
        if temp[r] < 1e-5:
            assert got[r] == ref == int(np.argmax(logits[r]))      # greedy: bit-exact index
        else:
            keep = osampler.top_p_keep_mask(logits[r] / temp[r], float(topp[r])) if topp[r] < 1 else np.ones(V, bool)
            assert keep[got[r]], r                                  # the draw is inside the reference's kept set
            n_same += int(got[r] == ref)
    assert n_same >= 7, n_same        # same counter RNG: draws agree except for ulp-level logf/expf ties (8 rows)


def test_sampler_distribution():
    """Draws over many steps follow softmax(logits / T) restricted to the top-p set."""
    rng = np.random.default_rng(5)
    V, n = 64, 4000
    l = (rng.standard_normal(V) * 1.5).astype(np.float32)
    logits = np.tile(l, (n, 1))
    temp = np.full(n, 0.8, np.float32)
    topp = np.full(n, 0.9, np.float32)
    seed = np.full(n, 12345, np.uint32)
    step = np.arange(n, dtype=np.int32)
    out = torch.empty(n, dtype=torch.int32, device="cuda")
    check(lib().b200_op_sample(ptr(dev(logits)), n, V, ptr(dev(temp)), ptr(dev(topp)), ptr(dev(seed.view(np.int32))),
                               ptr(dev(step)), ptr(out), stream()))
    got = out.cpu().numpy()
    keep = osampler.top_p_keep_mask(l / np.float32(0.8), 0.9)
    assert keep[got].all()
    p = np.exp((l / 0.8).astype(np.float64)) * keep
    p /= p.sum()
    freq = np.bincount(got, minlength=V) / n
    assert np.abs(freq - p).max() < 0.03

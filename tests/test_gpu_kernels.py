"""-m gpu: each C-ABI entry point against numpy / the oracle on seeded inputs (bit-exact for integer
work, stated tolerance for floating point)."""
import numpy as np
import pytest
import torch

import ctr_helpers as H

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hd():
    from deep_learning_b200 import _lib
    assert torch.cuda.is_available()
    return _lib.Handle(0)


def _st():
    return torch.cuda.current_stream().cuda_stream


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


@pytest.mark.parametrize("n,bits", [(1, 8), (31, 32), (4096, 16), (4097, 24), (100003, 27), (1 << 20, 32)])
def test_sort_pairs_stable_bit_exact(hd, n, bits):
    from deep_learning_b200._lib import ptr
    rng = np.random.RandomState(n)
    hi = (1 << bits) - 1
    keys = rng.randint(0, min(hi, 5000) + 1, size=n).astype(np.uint32) if n % 2 else rng.randint(0, hi + 1, size=n, dtype=np.int64).astype(np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    k = torch.from_numpy(keys.view(np.int32)).cuda(); v = torch.from_numpy(vals.view(np.int32)).cuda()
    wsn = hd.query("ctr_sort_ws_bytes", n)
    ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    hd.call("ctr_sort_pairs", _st(), ptr(k), ptr(v), n, bits, ptr(ws), wsn)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(_u32(k), keys[order])
    assert np.array_equal(_u32(v), vals[order])       # stability == bit-exact permutation


@pytest.mark.parametrize("n", [1, 2047, 2048, 2049, 70001])
def test_unique_segments_bit_exact(hd, n):
    from deep_learning_b200._lib import ptr
    rng = np.random.RandomState(n)
    keys = np.sort(rng.randint(0, max(2, n // 3), size=n).astype(np.uint32))
    k = torch.from_numpy(keys.view(np.int32)).cuda()
    uniq, uidx = torch.zeros(n, dtype=torch.int32, device="cuda"), torch.zeros(n, dtype=torch.int32, device="cuda")
    seg = torch.zeros(n + 2, dtype=torch.int32, device="cuda"); nu = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.zeros(4 * (n // 2048 + 8), dtype=torch.uint8, device="cuda")
    hd.call("ctr_unique_segments", _st(), ptr(k), n, ptr(uniq), ptr(seg), ptr(uidx), ptr(nu), ptr(ws), ws.numel(), None)
    ru, rstart, rinv = np.unique(keys, return_index=True, return_inverse=True)
    U = int(nu.item())
    assert U == len(ru)
    assert np.array_equal(_u32(uniq)[:U], ru)
    assert np.array_equal(_u32(seg)[:U + 1], np.r_[rstart, n].astype(np.uint32))
    assert np.array_equal(_u32(uidx), rinv.astype(np.uint32))


def test_sort_ids_sharded_keys_and_counts(hd):
    """owner-major key = (id % N) << bits | id // N; dropped zeros; shard counts (bit-exact shard map)"""
    from deep_learning_b200._lib import ptr
    rng = np.random.RandomState(7)
    B, F, N, bits = 300, 7, 4, 20
    ids = rng.randint(0, 5000, size=(B, F)).astype(np.int64)
    ids[rng.rand(B, F) < 0.2] = 0
    d = torch.from_numpy(ids).cuda()
    n = B * F + 3
    keys, perm = (torch.zeros(n, dtype=torch.int32, device="cuda") for _ in range(2))
    wsn = hd.query("ctr_sort_ws_bytes", n); ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    hd.call("ctr_sort_ids", _st(), ptr(d), 8, B, F, 0, F, 0, 1, 4000, 3, N, bits, 32, 0, None, ptr(keys), ptr(perm), ptr(ws), wsn)
    flat = np.r_[ids.reshape(-1), 4000 + np.arange(3)]
    pad = np.r_[ids.reshape(-1) == 0, np.zeros(3, bool)]
    ref = np.where(pad, 0xFFFFFFFF, ((flat % N) << bits) | (flat // N)).astype(np.uint32)
    order = np.argsort(ref, kind="stable")
    assert np.array_equal(_u32(keys), ref[order])
    assert np.array_equal(_u32(perm), np.arange(n, dtype=np.uint32)[order])
    uniq, uidx = torch.zeros(n, dtype=torch.int32, device="cuda"), torch.zeros(n, dtype=torch.int32, device="cuda")
    seg = torch.zeros(n + 2, dtype=torch.int32, device="cuda"); nu = torch.zeros(1, dtype=torch.int32, device="cuda")
    hd.call("ctr_unique_segments", _st(), ptr(keys), n, ptr(uniq), ptr(seg), ptr(uidx), ptr(nu), ptr(ws), wsn, None)
    cnt = torch.zeros(N, dtype=torch.int32, device="cuda")
    hd.call("ctr_shard_counts", _st(), ptr(uniq), ptr(nu), N, bits, ptr(cnt))
    ru = np.unique(ref); ru = ru[ru != 0xFFFFFFFF]
    assert np.array_equal(_u32(cnt), np.bincount(ru >> bits, minlength=N).astype(np.uint32))


@pytest.mark.parametrize("n,ties", [(2, False), (1000, True), (65536, False), (200001, True)])
def test_auc_matches_sklearn(hd, n, ties):
    from sklearn.metrics import roc_auc_score
    from deep_learning_b200._lib import ptr
    rng = np.random.RandomState(n)
    y = (rng.rand(n) < 0.3).astype(np.float32); y[0], y[1] = 1.0, 0.0
    s = rng.randn(n).astype(np.float32)
    if ties:
        s = np.round(s, 1); s[::7] = 0.0; s[3::11] = -0.0
    wsn = hd.query("ctr_auc_ws_bytes", n); ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    sd, yd = torch.from_numpy(s).cuda(), torch.from_numpy(y).cuda()   # keep both alive across the call
    hd.call("ctr_auc", _st(), ptr(sd), ptr(yd), n, ptr(out), ptr(ws), wsn)
    assert abs(out.item() - roc_auc_score(y, s)) < 1e-12


@pytest.mark.parametrize("B,K,N", [(1, 3, 5), (257, 170, 256), (1000, 429, 400), (300, 400, 1)])
def test_mlp_fp32_gemms(hd, B, K, N):
    """FP32 SIMT path vs fp64: forward (bias+relu), data gradient (mask) and weight gradient; 1e-5."""
    from deep_learning_b200._lib import ptr
    g = torch.Generator().manual_seed(B * 1000 + K)
    X = torch.randn(B, K, generator=g); W = torch.randn(K, N, generator=g) / K ** 0.5; b = torch.randn(N, generator=g)
    dZ = torch.randn(B, N, generator=g); mask = torch.randn(B, K, generator=g)
    Xc, Wc, bc, dZc, mc = (t.cuda() for t in (X, W, b, dZ, mask))
    Y = torch.zeros(B, N, device="cuda")
    hd.call("ctr_mlp_fwd", _st(), 0, B, K, N, ptr(Xc), K, ptr(Wc), None, ptr(bc), 1, 1.0, 0, ptr(Y), N)
    ref = torch.relu(X.double() @ W.double() + b.double())
    assert H.rel_err(Y.cpu(), ref) < 1e-5
    dX = torch.zeros(B, K, device="cuda")
    hd.call("ctr_mlp_bwd_data", _st(), 0, B, K, N, ptr(dZc), N, ptr(Wc), ptr(mc), K, 1.0, ptr(dX), K)
    ref = (dZ.double() @ W.double().T) * (mask > 0).double()
    assert H.rel_err(dX.cpu(), ref) < 1e-5
    wsn = hd.query("ctr_mlp_bwd_weight_ws_bytes", 0, B, K, N); ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    dW = torch.zeros(K, N, device="cuda"); db = torch.zeros(N, device="cuda")
    hd.call("ctr_mlp_bwd_weight", _st(), 0, B, K, N, ptr(Xc), K, ptr(dZc), N, ptr(dW), ptr(db), ptr(ws), wsn, 0)
    assert H.rel_err(dW.cpu(), X.double().T @ dZ.double()) < 1e-5
    assert H.rel_err(db.cpu(), dZ.double().sum(0)) < 1e-5


def test_table_pack_roundtrip_and_transpose(hd):
    from deep_learning_b200._lib import ptr
    R, k, stride = 1000, 16, 24
    emb = torch.randn(R, k, device="cuda"); first = torch.randn(R, device="cuda")
    tab = torch.full((R, stride), 7.0, device="cuda")
    hd.call("ctr_table_pack", _st(), ptr(emb), ptr(first), R, k, stride, ptr(tab))
    assert torch.equal(tab[:, :k], emb) and torch.equal(tab[:, k], first) and float(tab[:, k + 1:].abs().max()) == 0.0
    e2, f2 = torch.zeros_like(emb), torch.zeros_like(first)
    hd.call("ctr_table_unpack", _st(), ptr(tab), R, k, stride, ptr(e2), ptr(f2))
    assert torch.equal(e2, emb) and torch.equal(f2, first)
    W = torch.randn(429, 400, device="cuda"); Wt = torch.zeros(400, 429, device="cuda")
    hd.call("ctr_transpose", _st(), ptr(W), 429, 400, ptr(Wt), 429, 0)
    assert torch.equal(Wt, W.T.contiguous())


@pytest.mark.parametrize("B,K,N", [(128, 32, 16), (300, 170, 256), (1000, 429, 400), (4096, 400, 400), (2500, 256, 128),
                                   (777, 461, 512)])
def test_mlp_tf32_tcgen05_gemms(hd, B, K, N):
    """tcgen05/TMA TF32 path vs fp64: forward, data gradient and weight gradient within 2e-3."""
    from deep_learning_b200._lib import ptr
    if not hd.has_tcgen05:
        pytest.skip("needs sm_100")
    g = torch.Generator().manual_seed(B * 1000 + K)
    ldx = (K + 3) // 4 * 4
    X = torch.randn(B, K, generator=g); W = torch.randn(K, N, generator=g) / K ** 0.5; b = torch.randn(N, generator=g)
    dZ = torch.randn(B, N, generator=g); mask = torch.randn(B, K, generator=g)
    Xp = torch.zeros(B, ldx); Xp[:, :K] = X
    Xc, Wc, bc, dZc = Xp.cuda(), W.cuda(), b.cuda(), dZ.cuda()
    mp = torch.zeros(B, ldx); mp[:, :K] = mask; mc = mp.cuda()
    Wt = torch.zeros(N, ldx, device="cuda")
    hd.call("ctr_transpose", _st(), ptr(Wc), K, N, ptr(Wt), ldx, 1)
    Y = torch.zeros(B, N, device="cuda")
    hd.call("ctr_mlp_fwd", _st(), 1, B, K, N, ptr(Xc), ldx, ptr(Wc), ptr(Wt), ptr(bc), 1, 1.0, 0, ptr(Y), N)
    torch.cuda.synchronize()
    ref = torch.relu(X.double() @ W.double() + b.double())
    assert H.rel_err(Y.cpu(), ref) < 2e-3, "fwd"
    dX = torch.zeros(B, ldx, device="cuda")
    hd.call("ctr_mlp_bwd_data", _st(), 1, B, K, N, ptr(dZc), N, ptr(Wc), ptr(mc), ldx, 1.0, ptr(dX), ldx)
    torch.cuda.synchronize()
    ref = (dZ.double() @ W.double().T) * (mask > 0).double()
    assert H.rel_err(dX.cpu()[:, :K], ref) < 2e-3, "dgrad"
    wsn = hd.query("ctr_mlp_bwd_weight_ws_bytes", 1, B, K, N); ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    dW = torch.zeros(K, N, device="cuda"); db = torch.zeros(N, device="cuda")
    hd.call("ctr_mlp_bwd_weight", _st(), 1, B, K, N, ptr(Xc), ldx, ptr(dZc), N, ptr(dW), ptr(db), ptr(ws), wsn, 0)
    torch.cuda.synchronize()
    assert H.rel_err(dW.cpu(), X.double().T @ dZ.double()) < 2e-3, "wgrad"
    assert H.rel_err(db.cpu(), dZ.double().sum(0)) < 1e-5


@pytest.mark.parametrize("B,K,N", [(100, 37, 72), (1000, 429, 400), (4096, 400, 400), (1300, 400, 256), (777, 64, 500)])
def test_mlp_tf32_relu_sign_bits_equal_the_fp32_mask(hd, B, K, N):
    """ctr_mlp_fwd_bits / ctr_mlp_bwd_data_bits (ReLU sign bits in the kernels' own layout) against ctr_mlp_fwd /
    ctr_mlp_bwd_data with the fp32 activation as the mask: bit-for-bit the same Y and dX (tail tiles, both the
    one-tile and the persistent kernel, one and two MMA units)."""
    from deep_learning_b200._lib import ptr
    if not hd.has_tcgen05:
        pytest.skip("needs sm_100")
    g = torch.Generator().manual_seed(11)
    Kp = (K + 3) // 4 * 4
    X = torch.zeros(B, Kp)
    X[:, :K] = torch.randn(B, K, generator=g)
    X = X.cuda()
    W = (torch.randn(K, N, generator=g) * 0.1).cuda()
    b = torch.randn(N, generator=g).cuda()
    Wt = torch.zeros(N, Kp, device="cuda")
    hd.call("ctr_transpose", _st(), ptr(W), K, N, ptr(Wt), Kp, 1)
    ldy = (N + 1 + 31) // 32 * 32
    Y0, Y1 = torch.zeros(B, ldy, device="cuda"), torch.zeros(B, ldy, device="cuda")
    nb = hd.query("ctr_mlp_relu_bits_bytes", B, N)
    assert nb >= B * N // 8
    bits = torch.full((nb // 4,), -1, dtype=torch.int32, device="cuda")
    hd.call("ctr_mlp_fwd", _st(), 1, B, K, N, ptr(X), Kp, ptr(W), ptr(Wt), ptr(b), 1, 1.0, 0, ptr(Y0), ldy)
    hd.call("ctr_mlp_fwd_bits", _st(), B, K, N, ptr(X), Kp, ptr(Wt), ptr(b), 1.0, 0, ptr(Y1), ldy, ptr(bits))
    torch.cuda.synchronize()
    assert torch.equal(Y0, Y1)
    frac = (Y0[:, :N] > 0).float().mean().item()
    assert 0.2 < frac < 0.8                                   # both branches of the ReLU are exercised
    # the next layer's data gradient: dX[B, N] = (dZ[B, N2] @ W2[N, N2]^T) masked by Y
    N2 = 136
    dZ = torch.randn(B, N2, generator=g).cuda()
    W2 = (torch.randn(N, N2, generator=g) * 0.1).cuda()
    ldd = (N + 3) // 4 * 4
    d0, d1 = torch.zeros(B, ldd, device="cuda"), torch.zeros(B, ldd, device="cuda")
    hd.call("ctr_mlp_bwd_data", _st(), 1, B, N, N2, ptr(dZ), N2, ptr(W2), ptr(Y0), ldy, 1.0, ptr(d0), ldd)
    hd.call("ctr_mlp_bwd_data_bits", _st(), B, N, N2, ptr(dZ), N2, ptr(W2), ptr(bits), 1.0, ptr(d1), ldd)
    torch.cuda.synchronize()
    assert torch.equal(d0, d1)
    assert (d0[:, :N] != 0).float().mean().item() > 0.15


@pytest.mark.parametrize("B,H,n_fm,w_off,with_extra", [(1000, 400, 55, 55, False),     # config 4: w_deep slice 220 B into the head weight
                                                       (257, 128, 38, 38, False), (64, 512, 0, 0, True),
                                                       (77, 10, 7, 3, False),           # H % 4 != 0: scalar kernel
                                                       (4096, 400, 0, 20, True)])       # wdl: extra = wide sum
def test_head_fwd_both_kernels_match_fp64(hd, B, H, n_fm, w_off, with_extra):
    """ctr_head_fwd (models/deepfm.py:128-152 + the gradient through the last ReLU + head-weight gradients): the 128-bit
    kernel and the scalar one against an fp64 restatement, 1e-6 relative; pad columns of fm / deep hold garbage"""
    from deep_learning_b200._lib import ptr
    g = torch.Generator(device="cuda"); g.manual_seed(B + H)
    fm_ld, deep_ld, lddz = (n_fm + 3) // 4 * 4 + 4, H + 4 + (4 - H % 4) % 4, (H + 3) // 4 * 4
    fm = torch.randn(B, fm_ld, generator=g, device="cuda")
    deep = torch.relu(torch.randn(B, deep_ld, generator=g, device="cuda"))
    wall = torch.randn(w_off + H + 8, generator=g, device="cuda") * 0.1
    w_fm = torch.randn(max(n_fm, 1), generator=g, device="cuda") * 0.1
    w_deep = wall[w_off:w_off + H]
    bias = torch.randn(1, generator=g, device="cuda")
    extra = torch.randn(B, generator=g, device="cuda") if with_extra else None
    label = (torch.rand(B, generator=g, device="cuda") < 0.3).float()
    logit, score, dlogit = (torch.zeros(B, device="cuda") for _ in range(3))
    dz = torch.full((B, lddz), 7.0, device="cuda")
    loss = torch.zeros(1, device="cuda")
    g_fm, g_deep, g_bias = torch.zeros(max(n_fm, 1), device="cuda"), torch.zeros(H, device="cuda"), torch.zeros(1, device="cuda")
    wsn = hd.query("ctr_head_ws_bytes", B, n_fm, H)
    ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    hd.call("ctr_head_fwd", _st(), B, 2 * B, ptr(fm) if n_fm else None, fm_ld, n_fm, ptr(w_fm) if n_fm else None, ptr(deep), deep_ld, H,
            ptr(w_deep), ptr(extra), ptr(bias), ptr(label), ptr(logit), ptr(score), ptr(dlogit), ptr(dz), lddz, 0, ptr(loss),
            ptr(g_fm) if n_fm else None, ptr(g_deep), ptr(g_bias), ptr(ws), wsn, 0)
    d64 = lambda t: t.double().cpu()
    z = d64(deep[:, :H]) @ d64(w_deep) + d64(bias)
    if n_fm:
        z = z + d64(fm[:, :n_fm]) @ d64(w_fm[:n_fm])
    if with_extra:
        z = z + d64(extra)
    p = torch.sigmoid(z)
    y = d64(label)
    eps = 1e-7
    want_loss = (-y * torch.log(p + eps) - (1 - y) * torch.log(1 - p + eps)).sum() / (2 * B)
    dl = (-y / (p + eps) + (1 - y) / (1 - p + eps)) * p * (1 - p) / (2 * B)
    rel = lambda a, b: float((d64(a) - b).abs().max() / max(float(b.abs().max()), 1e-30))
    assert rel(logit, z) < 2e-6 and rel(score, p) < 2e-6 and rel(dlogit, dl) < 2e-5
    assert abs(float(loss) - float(want_loss)) < 2e-6 * max(1.0, abs(float(want_loss)))
    want_dz = dl[:, None] * d64(w_deep)[None, :] * (d64(deep[:, :H]) > 0)
    assert rel(dz[:, :H], want_dz) < 2e-5
    assert (dz[:, H:] == 7.0).all()                                       # nothing written past column H
    assert rel(g_deep, (dl[:, None] * d64(deep[:, :H])).sum(0)) < 2e-5
    assert abs(float(g_bias) - float(dl.sum())) < 2e-5 * float(dl.abs().sum())
    if n_fm:
        assert rel(g_fm[:n_fm], (dl[:, None] * d64(fm[:, :n_fm])).sum(0)) < 2e-5


def test_dense_update_equals_the_separate_launches(hd):
    """ctr_dense_update == ctr_adam_dense per regulariser group + ctr_transpose + ctr_round_tf32_copy, bit for bit"""
    import ctypes as C
    from deep_learning_b200._lib import ptr, AdamHP, DenseLayout
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    K, N = 37, 24
    ends = [1024, 2048, 3328]
    n = ends[2]
    th = torch.randn(n, generator=g, device="cuda"); th[100:110] = 0.0
    m, v = torch.randn(n, generator=g, device="cuda") * 0.01, torch.rand(n, generator=g, device="cuda") * 0.01
    gr = torch.randn(n, generator=g, device="cuda")
    hp = AdamHP(lr_t=1e-2, beta1=0.9, beta2=0.999, eps=1e-8)
    offs = [0, 1024 + 64, 2048 + 256]        # one [37, 24] weight matrix in each group, not at the group starts
    a = [t.clone() for t in (th, m, v)]
    for gi, (lo, hi) in enumerate(zip([0] + ends[:2], ends)):
        l2, l1 = (0.0, 0.0) if gi == 0 else ((0.3, 0.0) if gi == 1 else (0.0, 0.3))
        hd.call("ctr_adam_dense", _st(), ptr(a[0][lo:hi]), ptr(a[1][lo:hi]), ptr(a[2][lo:hi]), ptr(gr[lo:hi]), hi - lo, l2, l1,
                None, 1, hp)
    wt_a = [torch.zeros(N, 40, device="cuda") for _ in offs]
    wr_a = [torch.zeros(K, N, device="cuda") for _ in offs]
    for i, o in enumerate(offs):
        hd.call("ctr_transpose", _st(), ptr(a[0][o:o + K * N]), K, N, ptr(wt_a[i]), 40, 1)
        hd.call("ctr_round_tf32_copy", _st(), ptr(a[0][o:o + K * N]), ptr(wr_a[i]), K * N)
    b = [t.clone() for t in (th, m, v)]
    wt_b = [torch.zeros(N, 40, device="cuda") for _ in offs]
    wr_b = [torch.zeros(K, N, device="cuda") for _ in offs]
    lay = DenseLayout()
    for i in range(3):
        lay.group_end[i] = ends[i]
        lay.w_off[i], lay.K[i], lay.N[i], lay.ld_kmajor[i] = offs[i], K, N, 40
        lay.w_kmajor[i], lay.w_rounded[i] = wt_b[i].data_ptr(), wr_b[i].data_ptr()
    lay.reg_scale, lay.n_layers = 0.3, 3
    hd.call("ctr_dense_update", _st(), ptr(b[0]), ptr(b[1]), ptr(b[2]), ptr(gr), C.byref(lay), hp)
    torch.cuda.synchronize()
    for x, y in zip(a + wt_a + wr_a, b + wt_b + wr_b):
        assert torch.equal(x, y)

"""K4 entry points called directly through the C-ABI against numpy restatements (models/deepfm.py:78-98
differentiated by hand; TF1 Adam as tf.train.AdamOptimizer applies it to an IndexedSlices gradient with the
duplicates summed):

  ctr_embed_bwd_partials + ctr_embed_bwd_finalize   -> per-unique-row gradient [g_v | g_w | D]
  ctr_embed_bwd_partials + ctr_adam_rows (fused run sums, FM e-term, long runs on the warp-cooperative path)
  ctr_segment_sum_rows, ctr_scalar_segment_add

Sizes are chosen so that one hot row spans more than 8 chunks of 32 lookups (the long-run paths)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hd():
    from deep_learning_b200 import _lib
    return _lib.Handle(0)


def _ptr(t):
    return None if t is None else t.data_ptr()


def _st():
    return torch.cuda.current_stream().cuda_stream


def _u32(a):
    return torch.from_numpy(np.ascontiguousarray(a).astype(np.uint32).view(np.int32)).cuda()


def _sort_runs(hd, ids_t, B, Fc, rows, n_extra=0, extra_base=0):
    n = B * Fc + n_extra
    keys = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    perm = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    wsn = max(hd.query("ctr_sort_ws_bytes", n), 4 * (n // 2048 + 8)) + 256
    ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    hd.call("ctr_sort_ids", _st(), _ptr(ids_t), 4, B, Fc, 0, Fc, 0, 0, extra_base, n_extra, 1, 0,
            max(1, int(rows - 1).bit_length()), 0, None, _ptr(keys), _ptr(perm), _ptr(ws), wsn)
    uniq = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    seg = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    uidx = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    nun = torch.zeros(1, dtype=torch.int32, device="cuda")
    hd.call("ctr_unique_segments", _st(), _ptr(keys), n, _ptr(uniq), _ptr(seg), _ptr(uidx), _ptr(nun), _ptr(ws), wsn, None)
    return n, keys, perm, uniq, seg, uidx, nun


@pytest.mark.parametrize("k,Fc,C,B", [(16, 26, 13, 1500), (8, 20, 10, 700), (16, 3, 0, 4000), (4, 5, 2, 900)])
@pytest.mark.parametrize("use_fm", [1, 0])
def test_partials_finalize_and_row_adam(hd, k, Fc, C, B, use_fm):
    from deep_learning_b200._lib import AdamHP, EmbedCfg
    g = np.random.RandomState(7 + k + Fc)
    R = 3000
    has_first = use_fm
    stride = ((k + 1 + 7) // 8) * 8
    gs = ((k + 1 + 3) // 4) * 4
    cont_base = R - C
    n_extra = C if (use_fm and C) else 0
    ids = g.randint(0, cont_base, size=(B, Fc))
    ids[:, 0] = np.where(g.rand(B) < 0.6, 5, ids[:, 0])            # a hot row: ~0.6*B lookups -> > 8 chunks of 32
    x0_ld = ((Fc * k + C + 1 + 3) // 4) * 4
    fm_ld = ((Fc + C + k + 3) // 4) * 4
    dx0_ld = ((Fc * k + 3) // 4) * 4
    table = (g.standard_normal((R, stride)) * 0.1).astype(np.float32)
    m = (g.standard_normal((R, stride)) * 0.01).astype(np.float32)
    v = (g.rand(R, stride) * 0.01).astype(np.float32)
    dx0 = g.standard_normal((B, dx0_ld)).astype(np.float32)
    sum_e = g.standard_normal((B, k)).astype(np.float32)
    dlogit = (g.standard_normal(B) / B).astype(np.float32)
    w_fm = g.standard_normal(Fc + C + k).astype(np.float32)
    cont_grad = g.standard_normal((max(C, 1), gs)).astype(np.float32)
    x0 = np.zeros((B, x0_ld), np.float32)
    x0[:, :Fc * k] = table[:, :k][ids].reshape(B, Fc * k)

    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    ids_t = t(ids.astype(np.int32))
    n, keys, perm, uniq, seg, uidx, nun = _sort_runs(hd, ids_t, B, Fc, R, n_extra, cont_base)
    cfg = EmbedCfg(B=B, id_bytes=4, ids_ld=Fc, Fc=Fc, Fm=0, M=0, C=C, V=0, k=k, has_first=has_first, use_fm=use_fm,
                   fm_cont=use_fm, zero_row0=0, count_rule=0, combiner=0, table_stride=stride, x0_ld=x0_ld, fm_ld=fm_ld,
                   grad_stride=gs, round_tf32=0)
    part = torch.full(((n + n // 32 + 2) * gs,), float("nan"), device="cuda")
    t_dx0, t_x0, t_se, t_dl, t_w, t_cg = t(dx0), t(x0), t(sum_e), t(dlogit), t(w_fm), t(cont_grad)
    hd.call("ctr_embed_bwd_partials", _st(), cfg, _ptr(keys), _ptr(perm), _ptr(uidx), n, None, _ptr(t_x0), _ptr(t_dx0), dx0_ld,
            _ptr(t_se), _ptr(t_dl), _ptr(t_w), None, None, _ptr(t_cg), _ptr(part))

    # ---- numpy: per-row gradient [g_v | g_w | D] in the documented split form
    w2 = w_fm[Fc + C:Fc + C + k].astype(np.float64)
    G = np.zeros((R, k + 2), np.float64)
    lk = dx0[:, :Fc * k].reshape(B, Fc, k).astype(np.float64)
    if use_fm:
        lk = lk + dlogit[:, None, None].astype(np.float64) * w2[None, None, :] * sum_e[:, None, :].astype(np.float64)
    np.add.at(G[:, :k], ids, lk)
    if use_fm:
        np.add.at(G[:, k], ids, dlogit[:, None].astype(np.float64) * w_fm[None, :Fc].astype(np.float64))
        np.add.at(G[:, k + 1], ids, np.broadcast_to(dlogit[:, None].astype(np.float64), ids.shape))
        for c in range(n_extra):
            G[cont_base + c, :k] += cont_grad[c, :k]
            G[cont_base + c, k] += cont_grad[c, k]
    U = int(nun.item())
    rows_u = uniq[:U].cpu().numpy().astype(np.int64)
    assert np.array_equal(rows_u, np.unique(np.concatenate([ids.reshape(-1), cont_base + np.arange(n_extra)])))
    width = k + 2 if use_fm else k
    # ---- finalize (the multi-GPU path): run sums, hot row included
    grads = torch.full((U + 4, gs), float("nan"), device="cuda")
    hd.call("ctr_embed_bwd_finalize", _st(), _ptr(seg), _ptr(nun), n, width, gs, _ptr(part), _ptr(grads))
    got = grads[:U, :width].cpu().numpy().astype(np.float64)
    ref = G[rows_u][:, :width]
    scale = np.abs(ref).max(axis=0, keepdims=True) + 1e-30
    assert np.max(np.abs(got - ref) / scale) < 2e-5, "finalize"
    # ---- fused run sums + FM e-term + TF1 Adam on the touched rows
    hp = AdamHP(lr_t=0.01, beta1=0.9, beta2=0.999, eps=1e-8)
    t_tab, t_m, t_v = t(table), t(m), t(v)
    wsn = hd.query("ctr_adam_rows_ws_bytes", n)
    ws = torch.zeros(wsn, dtype=torch.uint8, device="cuda")
    touched = torch.zeros(R, dtype=torch.uint8, device="cuda")
    fm_w2 = t_w.data_ptr() + 4 * (Fc + C) if use_fm else None
    hd.call("ctr_adam_rows", _st(), _ptr(uniq), _ptr(nun), n, 0, _ptr(part), _ptr(seg), gs, k + has_first, _ptr(t_tab), _ptr(t_m),
            _ptr(t_v), stride, R, _ptr(touched), _ptr(ws), wsn, fm_w2, k, hp)
    torch.cuda.synchronize()
    gfull = G[:, :k + 1].copy()
    if use_fm:
        gfull[:, :k] -= w2[None, :] * table[:, :k].astype(np.float64) * G[:, k + 1:k + 2]       # the e-term
    wd = k + has_first
    gm = gfull[rows_u][:, :wd]
    m_ref = 0.9 * m[rows_u][:, :wd] + 0.1 * gm
    v_ref = 0.999 * v[rows_u][:, :wd] + 0.001 * gm * gm
    th_ref = table[rows_u][:, :wd] - 0.01 * m_ref / (np.sqrt(v_ref) + 1e-8)
    tab2, m2, v2 = t_tab.cpu().numpy(), t_m.cpu().numpy(), t_v.cpu().numpy()
    # fp32 sums of up to ~1000 addends (the hot row) against float64: a few 1e-6 relative on g, twice that on g^2
    def close(got, want, what):
        err = np.abs(got - want) - 5e-5 * np.abs(want)
        iu, ij = np.unravel_index(np.argmax(err), err.shape)
        assert err.max() < 1e-6, "%s: row %d col %d got %.7g want %.7g (run of %d lookups)" % (
            what, rows_u[iu], ij, got[iu, ij], want[iu, ij], int(seg[iu + 1].item()) - int(seg[iu].item()))
    close(m2[rows_u][:, :wd].astype(np.float64), m_ref, "Adam m")
    close(v2[rows_u][:, :wd].astype(np.float64), v_ref, "Adam v")
    close(tab2[rows_u][:, :wd].astype(np.float64), th_ref, "updated rows")
    mask = np.ones(R, bool)
    mask[rows_u] = False
    assert np.array_equal(tab2[mask], table[mask]) and np.array_equal(m2[mask], m[mask]), "untouched rows must not move"
    assert np.array_equal(touched.cpu().numpy().astype(bool), ~mask)


def test_segment_sum_rows_in_run_order(hd):
    g = np.random.RandomState(3)
    n, w, stride = 5000, 18, 20
    keys = np.sort(g.randint(0, 300, size=n)).astype(np.uint32)
    keys[100:1500] = keys[100]                                        # one long run
    keys = np.sort(keys)
    perm = g.permutation(n).astype(np.uint32)
    src = g.standard_normal((n, stride)).astype(np.float32)
    t_keys = _u32(keys)
    uniq = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    seg = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    uidx = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    nun = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.zeros(4 * (n // 2048 + 8) + 256, dtype=torch.uint8, device="cuda")
    hd.call("ctr_unique_segments", _st(), _ptr(t_keys), n, _ptr(uniq), _ptr(seg), _ptr(uidx), _ptr(nun), _ptr(ws), ws.numel(), None)
    U = int(nun.item())
    dst = torch.zeros(U, stride, device="cuda")
    t_src = torch.from_numpy(src).cuda()
    hd.call("ctr_segment_sum_rows", _st(), _ptr(seg), _ptr(nun), n, _ptr(_u32(perm)), _ptr(t_src), stride, w, _ptr(dst), stride)
    s = seg[:U + 1].cpu().numpy().astype(np.int64)
    ref = np.zeros((U, w), np.float32)
    for u in range(U):                    # sequential fp32 sum in run order: the kernel's order, so bit-exact
        acc = np.zeros(w, np.float32)
        for i in range(s[u], s[u + 1]):
            acc = acc + src[perm[i], :w]
        ref[u] = acc
    assert np.array_equal(dst[:, :w].cpu().numpy(), ref)


def test_scalar_segment_add_matches_add_at(hd):
    g = np.random.RandomState(5)
    B, ld, rows = 3000, 7, 500
    ids = g.randint(0, rows, size=(B, ld))
    ids[:, 1] = 3
    src = g.standard_normal(B).astype(np.float32)
    ids_t = torch.from_numpy(ids.astype(np.int32)).cuda()
    n, keys, perm, uniq, seg, uidx, nun = _sort_runs(hd, ids_t, B, ld, rows)
    dst0 = g.standard_normal(rows + 10).astype(np.float32)
    dst = torch.from_numpy(dst0.copy()).cuda()
    t_src = torch.from_numpy(src).cuda()
    hd.call("ctr_scalar_segment_add", _st(), _ptr(uniq), _ptr(seg), _ptr(nun), n, _ptr(perm), _ptr(t_src), ld, rows + 10, _ptr(dst),
            None, 0)
    ref = dst0.astype(np.float64)
    np.add.at(ref, ids, np.broadcast_to(src[:, None].astype(np.float64), ids.shape))
    got = dst.cpu().numpy()
    assert np.max(np.abs(got - ref)) < 1e-4 * max(1.0, np.abs(ref).max())
    assert np.array_equal(got[rows:], dst0[rows:])
    # with a workspace the run of 3000 (> 2048) is cut into pieces summed by one block each; twice in a row on the same
    # workspace (every call leaves its counters at zero), same result both times
    ws = torch.zeros(hd.query("ctr_scalar_segment_add_ws_bytes"), dtype=torch.uint8, device="cuda")
    outs = []
    for _ in range(2):
        d2 = torch.from_numpy(dst0.copy()).cuda()
        hd.call("ctr_scalar_segment_add", _st(), _ptr(uniq), _ptr(seg), _ptr(nun), n, _ptr(perm), _ptr(t_src), ld, rows + 10,
                _ptr(d2), _ptr(ws), ws.numel())
        outs.append(d2.cpu().numpy())
    assert np.array_equal(outs[0], outs[1])
    assert np.max(np.abs(outs[0] - ref)) < 1e-4 * max(1.0, np.abs(ref).max())
    assert int(ws[:8].view(torch.int32).abs().sum().item()) == 0


@pytest.mark.parametrize("k,has_first,use_fm", [(16, 1, 1), (16, 0, 0), (8, 1, 0), (32, 1, 1), (4, 0, 0)])
def test_adam_rows_gathered_equals_segment_sum_then_adam(hd, k, has_first, use_fm):
    """owner side of the sharded step: the fused sum-of-peer-contributions + TF1 Adam kernel equals
    ctr_segment_sum_rows followed by ctr_adam_rows bit for bit (same addition order), incl. masked keys (local_bits),
    a padding key and the FM e-term"""
    from deep_learning_b200._lib import AdamHP
    g = np.random.RandomState(100 + k + has_first)
    n, R, local_bits = 6000, 900, 10
    width = k + has_first
    wire = (k + 2 + 3) // 4 * 4 if use_fm else (width + 3) // 4 * 4
    stride = (width + 7) // 8 * 8
    rows = g.randint(0, R, size=n)
    owner_hi = g.randint(0, 4, size=n) << local_bits            # high bits = requesting peer's view: masked off
    keys = np.sort((rows | owner_hi).astype(np.uint32) & np.uint32((1 << local_bits) - 1)).astype(np.uint32)
    keys[-3:] = 0xFFFFFFFF                                       # padding run: skipped
    perm = g.permutation(n).astype(np.uint32)
    src = g.standard_normal((n, wire)).astype(np.float32)
    t_keys = _u32(keys)
    uniq = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    seg = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    uidx = torch.zeros(n + 8, dtype=torch.int32, device="cuda")
    nun = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = torch.zeros(4 * (n // 2048 + 8) + 256, dtype=torch.uint8, device="cuda")
    hd.call("ctr_unique_segments", _st(), _ptr(t_keys), n, _ptr(uniq), _ptr(seg), _ptr(uidx), _ptr(nun), _ptr(ws), ws.numel(), None)
    t_src, t_perm = torch.from_numpy(src).cuda(), _u32(perm)
    w2 = torch.from_numpy(g.standard_normal(k).astype(np.float32)).cuda() if use_fm else None
    hp = AdamHP(lr_t=1e-2, beta1=0.9, beta2=0.999, eps=1e-8)
    tabs = [torch.from_numpy(g.standard_normal((1 << local_bits, stride)).astype(np.float32)).cuda() for _ in range(2)]
    ms = [torch.from_numpy((g.standard_normal((1 << local_bits, stride)) * 0.01).astype(np.float32)).cuda() for _ in range(2)]
    vs = [torch.from_numpy((g.rand(1 << local_bits, stride) * 0.01).astype(np.float32)).cuda() for _ in range(2)]
    a = [tabs[0].clone(), ms[0].clone(), vs[0].clone()]
    b = [tabs[0].clone(), ms[0].clone(), vs[0].clone()]
    touched = [torch.zeros(1 << local_bits, dtype=torch.uint8, device="cuda") for _ in range(2)]
    # separate launches
    gw = k + 2 if use_fm else width
    own = torch.zeros(n, wire, device="cuda")
    hd.call("ctr_segment_sum_rows", _st(), _ptr(seg), _ptr(nun), n, _ptr(t_perm), _ptr(t_src), wire, gw, _ptr(own), wire)
    hd.call("ctr_adam_rows", _st(), _ptr(uniq), _ptr(nun), n, local_bits, _ptr(own), None, wire, width, _ptr(a[0]), _ptr(a[1]),
            _ptr(a[2]), stride, R, _ptr(touched[0]), None, 0, _ptr(w2), k, hp)
    # fused
    hd.call("ctr_adam_rows_gathered", _st(), _ptr(uniq), _ptr(nun), n, local_bits, _ptr(t_src), _ptr(t_perm), _ptr(seg), wire, width,
            _ptr(b[0]), _ptr(b[1]), _ptr(b[2]), stride, R, _ptr(touched[1]), _ptr(w2), k, hp)
    torch.cuda.synchronize()
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    assert torch.equal(touched[0], touched[1]) and int(touched[0].sum()) == len(np.unique(keys[:-3]))
    assert not torch.equal(a[0], tabs[0])

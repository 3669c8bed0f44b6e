"""K1 (ctr_fm_gather_fwd) called directly through the C-ABI, both kernels behind it -- the persistent fast path (no
pooled / vector block) and the general one (CTR_K1_GENERAL=1) -- against a numpy restatement of
models/deepfm.py:63-100: gathered rows and first-order terms bit-exact, the FM second-order term / sum_e within fp32
summation-order noise."""
import os

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


def tf32_rna(x):
    u = x.view(np.uint32).astype(np.uint64)
    return ((u + 0x1000) & 0xffffe000).astype(np.uint32).view(np.float32)


def k1_ref(table, ids, cont, cont_base, k, has_first, use_fm, fm_cont, zero0, x0_ld, fm_ld, rnd):
    """numpy restatement; table [R, stride] with columns 0..k-1 = embedding, column k = first-order weight"""
    B, Fc = ids.shape
    C = cont.shape[1]
    emb = table[:, :k][ids]                                  # [B, Fc, k]   models/deepfm.py:78
    first = table[:, k][ids] if has_first else np.zeros((B, Fc), np.float32)
    if zero0:
        dead = ids == 0
        emb = np.where(dead[..., None], 0.0, emb).astype(np.float32)
        first = np.where(dead, 0.0, first).astype(np.float32)
    x0 = np.zeros((B, x0_ld), np.float32)
    x0[:, :Fc * k] = emb.reshape(B, Fc * k)
    x0[:, Fc * k:Fc * k + C] = cont
    if rnd:
        x0 = tf32_rna(x0)
    x0[:, Fc * k + C] = 1.0
    if not use_fm:
        return x0, None, None
    fm = np.zeros((B, fm_ld), np.float32)
    fm[:, :Fc] = first
    e64 = emb.astype(np.float64)
    s, q = e64.sum(1), (e64 * e64).sum(1)
    if fm_cont and C:
        crow = table[cont_base:cont_base + C]
        ce = cont[:, :, None].astype(np.float32) * crow[None, :, :k]             # models/deepfm.py:85-87 (fp32 products)
        s, q = s + ce.astype(np.float64).sum(1), q + (ce.astype(np.float64) ** 2).sum(1)
        if has_first:
            fm[:, Fc:Fc + C] = crow[None, :, k] * cont
    fm[:, Fc + C:Fc + C + k] = 0.5 * (s * s - q)
    return x0, fm, s.astype(np.float32)


CASES = [  # k, Fc, C, id_bytes, use_fm, has_first, fm_cont, zero0, B
    (16, 26, 13, 4, 1, 1, 1, 0, 1000),     # config 4/5 shape
    (16, 26, 13, 8, 1, 1, 1, 0, 257),
    (8, 20, 10, 4, 1, 1, 1, 0, 256),       # config 1 shape
    (8, 20, 10, 8, 1, 1, 1, 1, 300),
    (16, 26, 13, 4, 0, 0, 0, 0, 513),      # DNN / WDL tower input only
    (16, 40, 5, 4, 1, 1, 1, 1, 100),       # > 32 fields: second id chunk
    (16, 1, 0, 4, 1, 1, 0, 0, 65),
    (4, 7, 3, 4, 1, 1, 1, 0, 129),
    (32, 9, 8, 4, 1, 1, 1, 1, 77),
    (32, 33, 0, 8, 1, 0, 0, 0, 31),
    (16, 26, 13, 4, 1, 1, 0, 0, 64),       # FM without the continuous rows
    (8, 5, 32, 4, 1, 1, 1, 0, 40),         # C = 32: no pad lanes left -> general kernel only
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("rnd", [0, 1])
@pytest.mark.parametrize("general", [0, 1])
def test_fm_gather_fwd_matches_restatement(hd, case, rnd, general):
    from deep_learning_b200._lib import EmbedCfg
    k, Fc, C, id_bytes, use_fm, has_first, fm_cont, zero0, B = case
    g = np.random.RandomState(1234 + k + Fc)
    R = 5000
    stride = ((k + 1 + 3) // 4) * 4 + (4 if k == 16 else 0)
    cont_base = R - C
    table = g.standard_normal((R, stride)).astype(np.float32)
    ids = g.randint(0, cont_base if cont_base > 0 else R, size=(B, Fc))
    ids[g.rand(B, Fc) < 0.15] = 0
    ids[:, 0] = g.randint(0, 4, size=B)                  # a hot field
    cont = g.rand(B, C).astype(np.float32)
    x0_ld = ((Fc * k + C + 1 + 3) // 4) * 4
    fm_ld = ((Fc + C + k + 3) // 4) * 4
    dev = "cuda"
    t_table = torch.from_numpy(table).to(dev)
    t_ids = torch.from_numpy(ids.astype(np.int64 if id_bytes == 8 else np.int32)).to(dev)
    t_cont = torch.from_numpy(cont).to(dev) if C else None
    t_x0 = torch.full((B, x0_ld), 7.0, device=dev)
    t_fm = torch.full((B, fm_ld), 7.0, device=dev)
    t_se = torch.full((B, k), 7.0, device=dev)
    cfg = EmbedCfg(B=B, id_bytes=id_bytes, ids_ld=Fc, Fc=Fc, Fm=0, M=0, C=C, V=0, k=k, has_first=has_first, use_fm=use_fm,
                   fm_cont=fm_cont, zero_row0=zero0, count_rule=0, combiner=0, table_stride=stride, x0_ld=x0_ld, fm_ld=fm_ld,
                   grad_stride=k + 4, round_tf32=rnd)
    crow = t_table[cont_base:] if (use_fm and fm_cont and C) else None
    old = os.environ.pop("CTR_K1_GENERAL", None)
    if general:
        os.environ["CTR_K1_GENERAL"] = "1"
    try:
        hd.call("ctr_fm_gather_fwd", torch.cuda.current_stream().cuda_stream, cfg, _ptr(t_table), _ptr(t_ids), 0, _ptr(t_cont),
                None, _ptr(crow), None, _ptr(t_x0), _ptr(t_fm), _ptr(t_se))
        torch.cuda.synchronize()
    finally:
        os.environ.pop("CTR_K1_GENERAL", None)
        if old is not None:
            os.environ["CTR_K1_GENERAL"] = old
    x0_ref, fm_ref, se_ref = k1_ref(table, ids, cont, cont_base, k, has_first, use_fm, fm_cont, zero0, x0_ld, fm_ld, rnd)
    D0 = Fc * k + C
    x0 = t_x0.cpu().numpy()
    assert np.array_equal(x0[:, :D0 + 1], x0_ref[:, :D0 + 1]), "X0 tile (bit-exact)"
    assert np.all(x0[:, D0 + 1:] == 0.0), "pad columns"
    if use_fm:
        fm = t_fm.cpu().numpy()
        assert np.array_equal(fm[:, :Fc], fm_ref[:, :Fc]), "first-order terms (bit-exact)"
        if C:
            assert np.array_equal(fm[:, Fc:Fc + C], fm_ref[:, Fc:Fc + C]), "continuous first-order terms"
        sec, sec_ref = fm[:, Fc + C:Fc + C + k], fm_ref[:, Fc + C:Fc + C + k]
        scale = max(1.0, float(np.abs(se_ref).max()) ** 2)
        assert np.max(np.abs(sec - sec_ref)) < 2e-6 * scale * Fc, "FM second-order"
        assert np.max(np.abs(t_se.cpu().numpy() - se_ref)) < 1e-5 * max(1.0, float(np.abs(se_ref).max())), "sum_e"

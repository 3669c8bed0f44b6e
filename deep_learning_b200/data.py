"""Synthetic Criteo-shaped batches and the reference's two batch formats.

gen-A (feed_dict generation, models/deepfm.py:173-180): `list[bytes]`, each
`pickle.dumps({cont_feats, vector_feats, cate_feats, wide_feats, labels, ...})` of Python lists
(utils/data_loader_load.py:130-137).
gen-B (tf.data generation, utils/data_loader.py:12-24): dict of arrays
`label [B,1] f32, cont_feats [B,C] f32, vector_feats [B,V] f32, cate_feats [B,Fc+M] i64`.
"""
from __future__ import annotations

import pickle
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

# cardinalities of the 26 categorical columns of the Criteo Kaggle display-ads set
CRITEO_CARDINALITIES = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27,
                        14992, 5461306, 10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]


def field_sizes(total_rows: int, n_fields: int) -> List[int]:
    """Split `total_rows` table rows over `n_fields` fields with Criteo-like proportions (>= 2 each)."""
    base = np.array([CRITEO_CARDINALITIES[i % 26] for i in range(n_fields)], dtype=np.float64)
    lo = 2 if total_rows >= 2 * n_fields else 1
    assert total_rows >= lo * n_fields, "need at least one row per field"
    spare = total_rows - lo * n_fields
    sz = lo + np.floor(base / base.sum() * spare).astype(np.int64)
    # hand the rounding remainder to the largest field so the ranges tile [0, total_rows) exactly
    sz[np.argmax(sz)] += total_rows - sz.sum()
    assert sz.min() >= 1 and sz.sum() == total_rows
    return [int(x) for x in sz]


def _draw_ids(gen: torch.Generator, shape, size: int, dist: str, alpha: float, device) -> torch.Tensor:
    """ids in [0, size): 'uniform', or bounded Zipf(alpha) ranks folded over the range by a
    multiplicative permutation so hot rows are scattered through the table."""
    if dist == "uniform" or size <= 2:
        return torch.randint(0, size, shape, generator=gen, device=device, dtype=torch.int64)
    u = torch.rand(shape, generator=gen, device=device, dtype=torch.float64)
    one_m_a = 1.0 - alpha
    rank = torch.floor(((size ** one_m_a - 1.0) * u + 1.0) ** (1.0 / one_m_a)).to(torch.int64).clamp_(1, size) - 1
    mult = 2654435761 % size
    while np.gcd(mult, size) != 1:
        mult += 1
    return (rank * mult) % size


def synthetic_batch(B: int, C: int, Fc: int, sizes: Sequence[int], seed: int, V: int = 0,
                    multi_top_n: Sequence[int] = (), multi_sizes: Sequence[int] = (), wide: int = 0,
                    dist: str = "zipf", alpha: float = 1.05, reserve_zero: bool = False,
                    id_dtype=torch.int32, device="cpu", pos_rate: float = 0.25) -> Dict[str, torch.Tensor]:
    """One gen-B style batch.  Field f draws from the id range [off_f, off_f + sizes[f]); with
    `reserve_zero` id 0 is kept for padding (row-0-zero models).  Multi-valued field j has
    `multi_top_n[j]` slots, length ~ U{0..top_n}, left-packed, 0-padded."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    cols, off = [], 1 if reserve_zero else 0
    all_sizes = list(sizes) + list(multi_sizes)
    assert len(sizes) == Fc and len(multi_sizes) == len(multi_top_n)
    for f in range(Fc):
        cols.append(_draw_ids(gen, (B, 1), all_sizes[f], dist, alpha, device) + off)
        off += all_sizes[f]
    for j, tn in enumerate(multi_top_n):
        ids = _draw_ids(gen, (B, tn), multi_sizes[j], dist, alpha, device) + off
        length = torch.randint(0, tn + 1, (B, 1), generator=gen, device=device)
        slot = torch.arange(tn, device=device).unsqueeze(0)
        cols.append(torch.where(slot < length, ids, torch.zeros_like(ids)))
        off += multi_sizes[j]
    cate = torch.cat(cols, dim=1).to(id_dtype) if cols else torch.zeros(B, 0, dtype=id_dtype, device=device)
    out = {
        "cate_feats": cate.contiguous(),
        "cont_feats": torch.rand(B, C, generator=gen, device=device, dtype=torch.float32),
        "vector_feats": torch.rand(B, V, generator=gen, device=device, dtype=torch.float32),
        "label": (torch.rand(B, 1, generator=gen, device=device) < pos_rate).to(torch.float32),
    }
    if wide:
        out["wide_feats"] = cate[:, :wide].contiguous()
    return out


def total_id_space(sizes: Sequence[int], multi_sizes: Sequence[int] = (), reserve_zero: bool = False) -> int:
    return int(sum(sizes) + sum(multi_sizes) + (1 if reserve_zero else 0))


def to_pickled_batch(batch: Dict[str, torch.Tensor]) -> bytes:
    """gen-A wire format: pickled dict of Python lists (utils/data_loader_load.py:130-137)."""
    d = {
        "cont_feats": batch["cont_feats"].cpu().tolist(),
        "vector_feats": batch["vector_feats"].cpu().tolist(),
        "cate_feats": batch["cate_feats"].cpu().tolist(),
        "mul_cate_feats": [], "mul_cate_feats_value": [],
        "wide_feats": batch["wide_feats"].cpu().tolist() if "wide_feats" in batch else [],
        "labels": batch["label"].cpu().tolist(),
    }
    return pickle.dumps(d)


def from_pickled_batch(blob: bytes, id_dtype=np.int32) -> Dict[str, np.ndarray]:
    """Decode a gen-A batch.  Values may be strings straight from str.split (utils/data_loader_load.py:24-25)."""
    d = pickle.loads(blob)
    n = len(d["cont_feats"])
    out = {
        "cont_feats": np.asarray(d["cont_feats"], dtype=np.float32).reshape(n, -1),
        "vector_feats": np.asarray(d.get("vector_feats", [[]] * n), dtype=np.float32).reshape(n, -1),
        "cate_feats": np.asarray(d["cate_feats"]).astype(id_dtype).reshape(n, -1),
        "label": np.asarray(d["labels"], dtype=np.float32).reshape(n, 1),
    }
    if len(d.get("wide_feats", [])) == n and n > 0:
        out["wide_feats"] = np.asarray(d["wide_feats"]).astype(id_dtype).reshape(n, -1)
    return out


def canonical_batch(d: Dict, spec) -> Dict:
    """Batch dict with the key names the engine reads.  The textline generation splits the ids over `cat_feats` /
    `cate_feats` + `mul_cat_feats` (+ `mul_cat_feats_value`) and calls the labels `labels`
    (utils/data_loader_textline.py:30-34, utils/data_loader_load.py:236-242, where the wide ids are stored as
    `wide_feats_value` while models/wdl_textline.py:62 reads `wide_feats`: both are accepted)."""
    if "cate_feats" in d and (spec.M == 0 or "mul_cat_feats" not in d) and "wide_feats_value" not in d:
        return d
    out = dict(d)
    single = d["cate_feats"] if "cate_feats" in d else d["cat_feats"]
    mul = d.get("mul_cat_feats", d.get("mul_cate_feats"))
    if spec.M and mul is not None and np.ndim(mul) == 2 and np.shape(single)[1] == spec.Fc:
        cat = torch.cat if isinstance(single, torch.Tensor) else np.concatenate
        mul = mul.to(single.dtype) if isinstance(single, torch.Tensor) else np.asarray(mul).astype(np.asarray(single).dtype)
        out["cate_feats"] = cat([single, mul], 1)
        val = d.get("mul_cat_feats_value", d.get("mul_cate_feats_value"))
        if val is not None:
            out["multi_feats_value"] = val
    else:
        out["cate_feats"] = single
    if "wide_feats" not in d and "wide_feats_value" in d:
        out["wide_feats"] = d["wide_feats_value"]
    if "label" not in d and "labels" in d:
        out["label"] = d["labels"]
    return out

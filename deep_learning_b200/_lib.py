"""ctypes binding of libctr_b200.so (the C-ABI declared in include/ctr_b200.h).

The product path has NO CPU fallback: if the library is missing or a call fails this module
raises; nothing is silently routed elsewhere.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libctr_b200.so")

c_i32, c_i64, c_u64, c_f32, c_vp = C.c_int32, C.c_int64, C.c_uint64, C.c_float, C.c_void_p


class EmbedCfg(C.Structure):
    """mirror of `ctr_embed_cfg` (include/ctr_b200.h)"""
    _fields_ = [(n, c_i32) for n in (
        "B", "id_bytes", "ids_ld", "Fc", "Fm", "M", "C", "V", "k", "has_first", "use_fm", "fm_cont",
        "zero_row0", "count_rule", "combiner", "table_stride", "x0_ld", "fm_ld", "grad_stride", "round_tf32",
        "reserved0")] + [("table_rows", c_i64), ("status", c_vp)]


HEAD_SCE_ON_SCORE, HEAD_NO_RELU_GATE = 1, 2          # CTR_HEAD_* flags of ctr_head_fwd
STATUS_WORDS = 4          # CTR_STATUS_WORDS: [step barrier, plan barrier, bad id, spare]


class AdamHP(C.Structure):
    """mirror of `ctr_adam_hp`"""
    _fields_ = [("lr_t", c_f32), ("beta1", c_f32), ("beta2", c_f32), ("eps", c_f32), ("c1", c_f32), ("c2", c_f32),
                ("guard", c_vp), ("kind", c_i32), ("reserved", c_i32)]
    ADAM, FTRL = 0, 1

    def __init__(self, lr_t, beta1=0.9, beta2=0.999, eps=1e-8, c1=None, c2=None, guard=None, kind=0):
        # c1 = 1 - beta1, c2 = 1 - beta2 from the exact (double) betas, rounded once (see include/ctr_b200.h);
        # guard: device pointer of the engine's status words (the update is skipped while they report a fault)
        super().__init__(lr_t, beta1, beta2, eps, 1.0 - beta1 if c1 is None else c1, 1.0 - beta2 if c2 is None else c2,
                         guard, kind, 0)


MAX_LAYERS = 8


class DenseLayout(C.Structure):
    """mirror of `ctr_dense_layout`"""
    _fields_ = [("group_end", c_i64 * 3), ("reg_scale", c_f32), ("n_layers", c_i32), ("w_off", c_i64 * MAX_LAYERS),
                ("K", c_i32 * MAX_LAYERS), ("N", c_i32 * MAX_LAYERS), ("ld_kmajor", c_i64 * MAX_LAYERS),
                ("w_kmajor", c_vp * MAX_LAYERS), ("w_rounded", c_vp * MAX_LAYERS)]


MAX_PEERS = 16
IPC_HANDLE_BYTES = 64


class PeerRoute(C.Structure):
    """mirror of `ctr_peer_route`"""
    _fields_ = [("world", c_i32), ("rank", c_i32), ("slab", c_vp * MAX_PEERS), ("seg_end", C.c_uint32 * MAX_PEERS),
                ("dst_row", C.c_uint32 * MAX_PEERS), ("dst_off", c_i64), ("dyn", c_vp)]


ROUTE_WORDS, PLAN_VEC, PLAN_GRAD, PLAN_NREQ, PLAN_U, PLAN_WORDS = 48, 0, 48, 96, 97, 128     # include/ctr_b200.h CTR_PLAN_*


class CrossCfg(C.Structure):
    """mirror of `ctr_cross_cfg`"""
    _fields_ = [("B", c_i32), ("D", c_i32), ("n_layers", c_i32), ("reserved", c_i32), ("ldx", c_i64),
                ("w", c_vp * MAX_LAYERS), ("b", c_vp * MAX_LAYERS), ("gw", c_vp * MAX_LAYERS), ("gb", c_vp * MAX_LAYERS)]


_P_CROSS = C.POINTER(CrossCfg)
_P_CFG = C.POINTER(EmbedCfg)
_P_RT = C.POINTER(PeerRoute)
_P_HP = C.POINTER(AdamHP)
_P_LAY = C.POINTER(DenseLayout)

# name -> (restype, [argtypes]) ; the first two args (handle, stream) are added for every entry
# except the handle-management ones.  Kept in header order.
_PROTOS = {
    "ctr_abi_version": (c_i32, None),
    "ctr_source_hash": (C.c_char_p, None),
    "ctr_sizeof_embed_cfg": (c_i32, None),
    "ctr_create": (c_i32, [c_i32, C.POINTER(c_vp)]),
    "ctr_destroy": (c_i32, [c_vp]),
    "ctr_last_error": (C.c_char_p, [c_vp]),
    "ctr_sm_count": (c_i32, [c_vp]),
    "ctr_launch_count": (c_i64, [c_vp]),
    "ctr_has_tcgen05": (c_i32, [c_vp]),
    "ctr_table_pack": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp]),
    "ctr_table_unpack": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp]),
    "ctr_pooled_gather_fwd": (c_i32, [c_vp, c_vp, _P_CFG, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "ctr_fm_gather_fwd": (c_i32, [c_vp, c_vp, _P_CFG, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "ctr_scalar_gather_sum": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i64, c_vp, c_vp]),
    "ctr_mlp_fwd": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp, c_i64, c_vp, c_vp, c_vp, c_i32, c_f32,
                            c_u64, c_vp, c_i64]),
    "ctr_mlp_bwd_data": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp, c_i64, c_vp, c_vp, c_i64, c_f32,
                                 c_vp, c_i64]),
    "ctr_mlp_relu_bits_bytes": (c_i64, [c_vp, c_i32, c_i32]),
    "ctr_mlp_fwd_bits": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_i64, c_vp, c_vp, c_f32, c_u64, c_vp, c_i64, c_vp]),
    "ctr_mlp_bwd_data_bits": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_i64, c_vp, c_vp, c_f32, c_vp, c_i64]),
    "ctr_mlp_bwd_weight_ws_bytes": (c_i64, [c_vp, c_i32, c_i32, c_i32, c_i32]),
    "ctr_mlp_bwd_weight": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp,
                                   c_vp, c_i64, c_i32]),
    "ctr_transpose": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_vp, c_i64, c_i32]),
    "ctr_round_tf32_copy": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64]),
    "ctr_cross_fwd": (c_i32, [c_vp, c_vp, _P_CROSS, c_vp, c_vp, c_i64]),
    "ctr_cross_bwd_ws_bytes": (c_i64, [c_vp, _P_CROSS]),
    "ctr_cross_bwd": (c_i32, [c_vp, c_vp, _P_CROSS, c_vp, c_vp, c_vp, c_vp, c_i64]),
    "ctr_head_ws_bytes": (c_i64, [c_vp, c_i32, c_i32, c_i32]),
    "ctr_head_fwd": (c_i32, [c_vp, c_vp, c_i32, c_i64, c_vp, c_i64, c_i32, c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_vp,
                             c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i32]),
    "ctr_head_bwd_deep": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp, c_i64, c_vp, c_i64]),
    "ctr_colsum_ws_bytes": (c_i64, [c_vp, c_i32, c_i32]),
    "ctr_colsum": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp, c_vp, c_i64]),
    "ctr_reduce": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_i64]),
    "ctr_reduce_static": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i32, c_vp]),
    "ctr_sort_ws_bytes": (c_i64, [c_vp, c_i64]),
    "ctr_sort_ids": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i64, c_i64, c_i32, c_i32, c_i64, c_i32, c_i64, c_i32,
                             c_i32, c_i32, c_i32, c_i64, c_vp, c_vp, c_vp, c_vp, c_i64]),
    "ctr_sort_pairs": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_vp, c_i64]),
    "ctr_unique_segments": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "ctr_shard_counts": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_vp]),
    "ctr_embed_bwd_partials": (c_i32, [c_vp, c_vp, _P_CFG, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_i64, c_vp,
                                       c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "ctr_embed_bwd_partials_weighted": (c_i32, [c_vp, c_vp, _P_CFG, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_i64, c_vp,
                                                c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "ctr_embed_bwd_finalize": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_vp]),
    "ctr_dense_update": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, _P_LAY, _P_HP]),
    "ctr_cont_bwd_ws_bytes": (c_i64, [c_vp, _P_CFG]),
    "ctr_fm_cont_bwd": (c_i32, [c_vp, c_vp, _P_CFG, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64]),
    "ctr_scalar_segment_add": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_i32, c_i64, c_vp, c_vp, c_i64]),
    "ctr_scalar_segment_add_ws_bytes": (c_i64, [c_vp]),
    "ctr_scalar_gathered_l2": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_f32, c_i64, c_i64, c_i32, c_f32, c_vp,
                                       c_vp]),
    "ctr_segment_sum_rows": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_i32, c_i32, c_vp, c_i32]),
    "ctr_adam_rows_ws_bytes": (c_i64, [c_vp, c_i64]),
    "ctr_adam_rows": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp,
                              c_i32, c_i64, c_vp, c_vp, c_i64, c_vp, c_i32, _P_HP]),
    "ctr_adam_rows_gathered": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp,
                                       c_i32, c_i64, c_vp, c_vp, c_i32, _P_HP]),
    "ctr_adam_dense": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_f32, c_f32, c_vp, c_i32, _P_HP]),
    "ctr_auc_ws_bytes": (c_i64, [c_vp, c_i64]),
    "ctr_auc": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_i64]),
    "ctr_gather_rows": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_i32]),
    "ctr_peer_alloc": (c_i32, [c_vp, c_i64, C.POINTER(c_vp), c_vp]),
    "ctr_peer_open": (c_i32, [c_vp, c_vp, C.POINTER(c_vp)]),
    "ctr_peer_close": (c_i32, [c_vp, c_vp]),
    "ctr_peer_free": (c_i32, [c_vp, c_vp]),
    "ctr_peer_barrier": (c_i32, [c_vp, c_vp, _P_RT, c_i64, C.c_uint32, c_vp, c_i32]),
    "ctr_peer_bcast": (c_i32, [c_vp, c_vp, c_vp, c_i32, _P_RT]),
    "ctr_peer_push_rows": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i32, _P_RT]),
    "ctr_peer_reduce_sum": (c_i32, [c_vp, c_vp, _P_RT, c_i64, c_vp]),
    "ctr_gather_rows_push": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_vp, c_i64, c_i32, c_i32, c_i32, _P_RT]),
    "ctr_embed_bwd_finalize_push": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, _P_RT]),
    "ctr_merge_sorted_segments": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "ctr_peer_plan_routes": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i64, c_i64, c_vp, c_vp]),
    "ctr_zero_row_at": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_vp]),
    "ctr_debug_gemm_trace": (c_i32, [c_vp, c_i32]),
    "ctr_invert_perm": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "ctr_parse_lines": (c_i32, [c_vp, c_i64, c_i32, c_i32, c_i32, c_i32, c_i32, c_i64, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp,
                                c_vp, c_vp, c_vp]),
    "ctr_parse_tfrecords": (c_i32, [c_vp, c_i64, c_i32, c_i32, c_i32, c_i32, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
}

EXPORTS = tuple(_PROTOS)


class CtrError(RuntimeError):
    pass


_lib = None


def load(path: str = LIB_PATH) -> C.CDLL:
    """dlopen the C-ABI library and set prototypes.  Raises if it is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise CtrError("libctr_b200.so not found at %s; build it with `python -m deep_learning_b200.build` "
                       "(there is no CPU fallback)" % path)
    lib = C.CDLL(path)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        if args is not None:
            fn.argtypes = args
    if lib.ctr_abi_version() != 4:
        raise CtrError("ABI version mismatch")
    if lib.ctr_sizeof_embed_cfg() != C.sizeof(EmbedCfg):
        raise CtrError("ctr_embed_cfg layout mismatch")
    # the binary must be the one these sources build (a stale .so with fresh timestamps would otherwise go unnoticed);
    # CTR_ALLOW_STALE_LIB=1 for a deployment that ships the library without csrc/
    if path == LIB_PATH and os.environ.get("CTR_ALLOW_STALE_LIB") != "1":
        from . import build as _build
        if os.path.isdir(_build.CSRC):
            want, have = _build.source_hash(), lib.ctr_source_hash().decode()
            if want != have:
                raise CtrError("libctr_b200.so was built from other sources (library %s..., sources %s...): rebuild with "
                               "`python -m deep_learning_b200.build`" % (have[:12], want[:12]))
    _lib = lib
    return lib


def ptr(t):
    """device pointer of a torch tensor (None -> NULL)"""
    return None if t is None else t.data_ptr()


class Handle:
    """one `ctr_handle` per device; `call` checks the status code and raises with ctr_last_error."""

    def __init__(self, device_index: int):
        self.lib = load()
        h = c_vp()
        rc = self.lib.ctr_create(int(device_index), C.byref(h))
        if rc != 0 or not h.value:
            raise CtrError("ctr_create(device=%d) failed with %d (is a CUDA device present?)" % (device_index, rc))
        self.h = h
        self.device_index = device_index
        self.sm_count = self.lib.ctr_sm_count(h)
        self.has_tcgen05 = bool(self.lib.ctr_has_tcgen05(h))
        self.timers = None     # name -> [events] when bench.py profiles selected entry points
        self._fn = {}          # entry point name -> bound ctypes function (getattr on a CDLL is a dict miss + a dlsym wrapper)

    def call(self, name: str, stream: int, *args):
        tm = self.timers.get(name) if self.timers is not None else None
        if tm is not None:
            import torch
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ts = torch.cuda.ExternalStream(stream) if stream else torch.cuda.current_stream()
            e0.record(ts)        # events go on the stream the kernel is launched on
        fn = self._fn.get(name)
        if fn is None:
            fn = self._fn[name] = getattr(self.lib, name)
        rc = fn(self.h, stream, *args)
        if rc != 0:
            raise CtrError("%s failed (%d): %s" % (name, rc, self.lib.ctr_last_error(self.h).decode()))
        if tm is not None:
            e1.record(ts)
            tm.append((e0, e1))

    @property
    def launches(self) -> int:
        """kernels launched through this handle (counted inside the library)"""
        return int(self.lib.ctr_launch_count(self.h))

    def invoke(self, name: str, *args) -> None:
        """a status-returning entry point without a stream argument (set-up calls)"""
        rc = getattr(self.lib, name)(self.h, *args)
        if rc != 0:
            raise CtrError("%s failed (%d): %s" % (name, rc, self.lib.ctr_last_error(self.h).decode()))

    def query(self, name: str, *args) -> int:
        return int(getattr(self.lib, name)(self.h, *args))

    def close(self):
        if self.h is not None and self.h.value:
            self.lib.ctr_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

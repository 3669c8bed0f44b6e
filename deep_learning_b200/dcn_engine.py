"""DcnEngine: the train / predict step of the reference's test_model/DCN.py on one GPU.

The model has no embedding table: a dense feature row `x0` [B, D] feeds (a) a tower of LINEAR layers (matmul + bias, no
activation: test_model/DCN.py:66-70) and (b) a cross network (:73-88); `[cross_out | deep_out] @ concat_projection +
concat_bias` -> sigmoid -> tf.losses.log_loss + l2 on concat_projection (:90-111); tf.train.AdamOptimizer (:116).
Every arithmetic step is a call into libctr_b200.so: ctr_mlp_fwd / ctr_mlp_bwd_* (relu off), ctr_cross_fwd / ctr_cross_bwd,
ctr_head_fwd (CTR_HEAD_NO_RELU_GATE), ctr_reduce, ctr_dense_update.  FP32 tower (the tcgen05 kernels want shapes an
arbitrary feature width does not have).

Reference variables <-> flat buffer: `concat_projection` [D+H, 1] is kept as two separately aligned halves (cross part,
deep part); everything else by its own name.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, List, Optional

import numpy as np
import torch

from . import _lib
from ._lib import AdamHP, CrossCfg, DenseLayout, ptr
from .engine import ADAM_BETA1, ADAM_BETA2, ADAM_EPS, _ALIGN, _ru

CROSS_LAYERS = 2          # test_model/DCN.py:28


class DcnEngine:
    def __init__(self, deep_feature_size: int, hidden_units: List[int], max_batch: int, device=None,
                 learning_rate: float = 0.001, decay_steps: int = 10000000, decay_rate: float = 0.9, l2_reg: float = 0.0,
                 cross_layers: int = CROSS_LAYERS, train: bool = True):
        if not torch.cuda.is_available():
            raise _lib.CtrError("DcnEngine needs a CUDA device: the hot path has no CPU fallback")
        self.D, self.hidden, self.Bmax, self.Lc = int(deep_feature_size), [int(h) for h in hidden_units], int(max_batch), int(cross_layers)
        self.dev = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.h = _lib.Handle(self.dev.index if self.dev.index is not None else torch.cuda.current_device())
        self.lr0, self.decay_steps, self.decay_rate, self.l2_reg = learning_rate, decay_steps, decay_rate, float(l2_reg)
        self.train_mode = train
        self.global_step, self.beta1_power, self.beta2_power = 0, ADAM_BETA1, ADAM_BETA2
        D, H, B = self.D, self.hidden[-1], self.Bmax
        f32 = dict(dtype=torch.float32, device=self.dev)
        # ---- flat parameter buffer: [no regulariser ... | l2: the two halves of concat_projection]
        shapes = []
        prev = D
        for i, h_ in enumerate(self.hidden):
            shapes += [("deep_layer_%d" % i, (prev, h_)), ("deep_layer_bias_%d" % i, (1, h_))]
            prev = h_
        for i in range(self.Lc):
            shapes += [("cross_layer_%d" % i, (D, 1)), ("cross_bias_%d" % i, (D, 1))]
        shapes += [("concat_bias", (1,))]
        self.off: Dict[str, int] = {}
        self.shape = dict(shapes)
        cur = 0
        for n, shp in shapes:
            self.off[n] = cur
            cur += _ru(int(np.prod(shp)), _ALIGN)
        self.l2_start = cur
        self.off["proj_cross"], self.shape["proj_cross"] = cur, (D,)
        cur += _ru(D, _ALIGN)
        self.off["proj_deep"], self.shape["proj_deep"] = cur, (H,)
        cur += _ru(H, _ALIGN)
        self.n_dense = cur
        self.theta = torch.zeros(cur, **f32)
        # ---- activations
        self.x0 = torch.zeros(B, _ru(D, 4), **f32)
        self.acts = [torch.zeros(B, _ru(h_, 4), **f32) for h_ in self.hidden]
        self.cross_out = torch.zeros(B, _ru(D, 4), **f32)
        self.label = torch.zeros(B, **f32)
        self.logit, self.score, self.dlogit = (torch.zeros(B, **f32) for _ in range(3))
        self.loss, self.reg_term = torch.zeros(1, **f32), torch.zeros(1, **f32)
        ws_bytes = max(self.h.query("ctr_head_ws_bytes", B, D, H), 1 << 16)
        if train:
            self.m, self.v = torch.zeros(cur, **f32), torch.zeros(cur, **f32)
            self.grad = torch.zeros(cur, **f32)
            self.dz = [torch.zeros(B, max(self.hidden), **f32) for _ in range(2)]
            prev = D
            for h_ in self.hidden:
                ws_bytes = max(ws_bytes, self.h.query("ctr_mlp_bwd_weight_ws_bytes", 0, B, prev, h_))
                prev = h_
            ws_bytes = max(ws_bytes, self.h.query("ctr_cross_bwd_ws_bytes", C.byref(self._cross_cfg(B, with_grads=True))))
            self.status = torch.zeros(_lib.STATUS_WORDS, dtype=torch.int32, device=self.dev)
        self.ws = torch.zeros(_ru(ws_bytes, 256), dtype=torch.uint8, device=self.dev)
        self.last_batch = 0
        self._lay = None

    # ---- views ------------------------------------------------------------------------------------------------
    def view(self, name: str, buf: Optional[torch.Tensor] = None) -> torch.Tensor:
        buf = self.theta if buf is None else buf
        return buf[self.off[name]:self.off[name] + int(np.prod(self.shape[name]))]

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.dev).cuda_stream

    def _cross_cfg(self, B: int, with_grads: bool = False) -> CrossCfg:
        c = CrossCfg()
        c.B, c.D, c.n_layers, c.reserved, c.ldx = B, self.D, self.Lc, 0, self.x0.shape[1]
        for l in range(self.Lc):
            c.w[l], c.b[l] = self.view("cross_layer_%d" % l).data_ptr(), self.view("cross_bias_%d" % l).data_ptr()
            if with_grads:
                c.gw[l] = self.view("cross_layer_%d" % l, self.grad).data_ptr()
                c.gb[l] = self.view("cross_bias_%d" % l, self.grad).data_ptr()
        return c

    # ---- parameters: reference names and shapes (test_model/DCN.py:49-99) ----------------------------------------
    def _names(self):
        return [n for n in self.off if n not in ("proj_cross", "proj_deep")] + ["concat_projection"]

    def _load(self, p: Dict[str, np.ndarray], flat: torch.Tensor) -> None:
        t = lambda a: torch.as_tensor(np.ascontiguousarray(np.asarray(a, dtype=np.float32))).to(self.dev)
        for n in self.off:
            if n in ("proj_cross", "proj_deep"):
                continue
            self.view(n, flat).copy_(t(p[n]).reshape(-1))
        proj = t(p["concat_projection"]).reshape(-1)
        self.view("proj_cross", flat).copy_(proj[:self.D])
        self.view("proj_deep", flat).copy_(proj[self.D:])

    def _dump(self, flat: torch.Tensor) -> Dict[str, np.ndarray]:
        out = {}
        for n in self.off:
            if n in ("proj_cross", "proj_deep"):
                continue
            shp = self.shape[n] if n != "concat_bias" else ()
            out[n] = self.view(n, flat).cpu().numpy().reshape(shp).copy()
        out["concat_projection"] = torch.cat([self.view("proj_cross", flat), self.view("proj_deep", flat)]).cpu().numpy().reshape(-1, 1)
        return out

    def set_weights(self, p: Dict[str, np.ndarray]) -> None:
        self._load(p, self.theta)

    def get_weights(self) -> Dict[str, np.ndarray]:
        return self._dump(self.theta)

    def set_optimizer_state(self, m, v, global_step: int, beta1_power: float, beta2_power: float) -> None:
        self._load(m, self.m)
        self._load(v, self.v)
        self.global_step, self.beta1_power, self.beta2_power = int(global_step), float(beta1_power), float(beta2_power)

    def get_optimizer_state(self):
        return self._dump(self.m), self._dump(self.v)

    def check_status(self, wait: bool = True) -> None:
        """no ids, no peers: nothing can set a status word here (kept for the shared fit loops)"""

    # ---- step ----------------------------------------------------------------------------------------------------
    def load_batch(self, batch: Dict) -> int:
        x = batch["features"] if "features" in batch else batch["feature_values"]
        x = x if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x, dtype=np.float32))
        B = int(x.shape[0])
        if B > self.Bmax or x.shape[1] != self.D:
            raise ValueError("features is [%d, %d], engine takes up to [%d, %d]" % (B, x.shape[1], self.Bmax, self.D))
        self.x0[:B, :self.D].copy_(x.to(torch.float32), non_blocking=True)
        lab = batch.get("label", batch.get("labels"))
        if lab is not None:
            lab = lab if isinstance(lab, torch.Tensor) else torch.as_tensor(np.asarray(lab, dtype=np.float32))
            self.label[:B].copy_(lab.reshape(-1).to(torch.float32), non_blocking=True)
        self.last_batch = B
        return B

    def forward(self, B: int, with_loss: bool) -> None:
        st, h, D, H = self._stream(), self.h, self.D, self.hidden[-1]
        x, ldx, k_ = self.x0, self.x0.shape[1], D
        for i, h_ in enumerate(self.hidden):                 # test_model/DCN.py:66-70: matmul + bias, NO activation
            h.call("ctr_mlp_fwd", st, 0, B, k_, h_, ptr(x), ldx, ptr(self.view("deep_layer_%d" % i)), None,
                   ptr(self.view("deep_layer_bias_%d" % i)), 0, 1.0, 0, ptr(self.acts[i]), self.acts[i].shape[1])
            x, ldx, k_ = self.acts[i], self.acts[i].shape[1], h_
        h.call("ctr_cross_fwd", st, C.byref(self._cross_cfg(B)), ptr(self.x0), ptr(self.cross_out), self.cross_out.shape[1])
        train = with_loss and self.train_mode
        g = self.grad if train else None
        h.call("ctr_head_fwd", st, B, B, ptr(self.cross_out), self.cross_out.shape[1], D, ptr(self.view("proj_cross")), ptr(x), ldx, H,
               ptr(self.view("proj_deep")), None, ptr(self.view("concat_bias")), ptr(self.label) if with_loss else None,
               ptr(self.logit), ptr(self.score), ptr(self.dlogit) if with_loss else None,
               ptr(self.dz[0]) if train else None, H, 0, ptr(self.loss) if with_loss else None,
               ptr(self.view("proj_cross", g)) if train else None, ptr(self.view("proj_deep", g)) if train else None,
               ptr(self.view("concat_bias", g)) if train else None, ptr(self.ws), self.ws.numel(), _lib.HEAD_NO_RELU_GATE)

    def _adam_hp(self) -> AdamHP:
        lr = self.lr0 * self.decay_rate ** math.floor(self.global_step / self.decay_steps)
        lr_t = lr * math.sqrt(1.0 - self.beta2_power) / (1.0 - self.beta1_power)
        return AdamHP(lr_t=lr_t, beta1=ADAM_BETA1, beta2=ADAM_BETA2, eps=ADAM_EPS, guard=self.status.data_ptr())

    def _layout(self) -> DenseLayout:
        if self._lay is None:
            lay = DenseLayout()
            lay.group_end[0], lay.group_end[1], lay.group_end[2] = self.l2_start, self.n_dense, self.n_dense
            lay.reg_scale, lay.n_layers = self.l2_reg, 0
            self._lay = lay
        return self._lay

    def train_step(self, batch: Optional[Dict] = None, next_batch: Optional[Dict] = None) -> torch.Tensor:
        """One `sess.run([loss, optimizer])` (test_model/DCN.py:136).  Returns the loss (log-loss + l2 term) as a device scalar."""
        if not self.train_mode:
            raise _lib.CtrError("engine was created with train=False")
        st, h, D = self._stream(), self.h, self.D
        B = self.load_batch(batch) if batch is not None else self.last_batch
        self.forward(B, with_loss=True)
        reg = None
        if self.l2_reg > 0:
            h.call("ctr_reduce", st, ptr(self.theta[self.l2_start:]), self.n_dense - self.l2_start, 1, ptr(self.reg_term), ptr(self.ws), self.ws.numel())
            reg = self.reg_term * (0.5 * self.l2_reg)
        loss = self.loss.clone()
        # ---- backward: cross network (needs dlogit only), then the linear tower
        h.call("ctr_cross_bwd", st, C.byref(self._cross_cfg(B, with_grads=True)), ptr(self.x0), ptr(self.dlogit),
               ptr(self.view("proj_cross")), ptr(self.ws), self.ws.numel())
        cur = 0
        dims = [(self.D if i == 0 else self.hidden[i - 1], h_) for i, h_ in enumerate(self.hidden)]
        for i in range(len(self.hidden) - 1, -1, -1):
            k_, h_ = dims[i]
            x, ldx = (self.acts[i - 1], self.acts[i - 1].shape[1]) if i > 0 else (self.x0, self.x0.shape[1])
            h.call("ctr_mlp_bwd_weight", st, 0, B, k_, h_, ptr(x), ldx, ptr(self.dz[cur]), h_, ptr(self.view("deep_layer_%d" % i, self.grad)),
                   ptr(self.view("deep_layer_bias_%d" % i, self.grad)), ptr(self.ws), self.ws.numel(), 0)
            if i > 0:            # no activation: no mask
                h.call("ctr_mlp_bwd_data", st, 0, B, k_, h_, ptr(self.dz[cur]), h_, ptr(self.view("deep_layer_%d" % i)), None, 0, 1.0,
                       ptr(self.dz[cur ^ 1]), k_)
                cur ^= 1
        h.call("ctr_dense_update", st, ptr(self.theta), ptr(self.m), ptr(self.v), ptr(self.grad), C.byref(self._layout()), self._adam_hp())
        self.beta1_power *= ADAM_BETA1
        self.beta2_power *= ADAM_BETA2
        self.global_step += 1
        return loss if reg is None else loss + reg

    def predict(self, batch: Optional[Dict] = None) -> torch.Tensor:
        """`session.run("Sigmoid:0")` (test_model/DCN.py:190): scores [B] on the device."""
        B = self.load_batch(batch) if batch is not None else self.last_batch
        self.forward(B, with_loss=False)
        return self.score[:B]

    def auc(self, scores: torch.Tensor, labels: torch.Tensor) -> float:
        n = int(scores.numel())
        sc = scores.reshape(-1).to(self.dev, torch.float32).contiguous()
        lb = labels.reshape(-1).to(self.dev, torch.float32).contiguous()
        wsn = self.h.query("ctr_auc_ws_bytes", n)
        ws = torch.empty(wsn, dtype=torch.uint8, device=self.dev)
        out = torch.zeros(1, dtype=torch.float64, device=self.dev)
        self.h.call("ctr_auc", self._stream(), ptr(sc), ptr(lb), n, ptr(out), ptr(ws), wsn)
        return float(out.item())

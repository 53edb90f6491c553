"""Host-side mirror of the reference's ``li.model`` for INFERENCE (search/li/model.py).

``NeuralNetwork`` keeps the reference's public surface on the query path --
``predict_proba(X) -> (probs f32[nq,C] descending, classes i64[nq,C])`` (:226-241) and
``predict(X) -> argmax`` (:213-224) -- but the arithmetic runs in the sm_100a kernels
behind ``lmi_linear`` / ``lmi_softmax_topn`` / ``lmi_argmax``.  Training
(``train`` / ``train_batch``, :166-211) is the offline build and out of scope.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native

# hidden widths of ``supported_models`` -- search/li/model.py:29-79
HIDDEN_LAYERS = {
    "MLP": [128], "MLP-2": [64], "MLP-3": [256], "MLP-4": [512], "MLP-5": [256, 128],
    "MLP-6": [32], "MLP-7": [16], "MLP-8": [8], "MLP-9": [8, 16],
}


# lmi_mlp_topn keeps per-(query tile, hidden chunk) partial logits in its workspace (4 KB per query for MLP-4 with
# 120 buckets): fine for a query batch, too much for the million-row chunks of the layout build's predict()
FUSED_LOGITS_MAX_ROWS = 32768


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def data_X_to_torch(data) -> torch.Tensor:
    """search/li/model.py:102-105."""
    return torch.from_numpy(np.array(data).astype(np.float32))


def get_device() -> torch.device:
    """search/li/model.py:115-127 -- but there is no CPU path here: CUDA is required."""
    if not torch.cuda.is_available():
        raise RuntimeError("learnedmetricindex_b200 needs a CUDA device (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def to_device_matrix(x, device: torch.device) -> torch.Tensor:
    """Host ndarray / torch tensor -> contiguous device matrix, fp16 kept as fp16 (widening
    to f32 happens exactly inside the kernels), everything else as f32 (model.py:104)."""
    if isinstance(x, torch.Tensor):
        t = x
    else:
        a = np.asarray(x)
        if a.dtype not in (np.float16, np.float32):
            a = a.astype(np.float32)
        t = torch.from_numpy(np.ascontiguousarray(a))
    if t.dtype not in (torch.float16, torch.float32):
        t = t.float()
    return t.to(device, non_blocking=True).contiguous()


class NeuralNetwork:
    """Inference-only mirror of ``li.model.NeuralNetwork`` holding the weights on the GPU."""

    def __init__(self, layers: Sequence[Tuple[object, object]], model_type: Optional[str] = None,
                 device: Optional[torch.device] = None):
        self.device = device or get_device()
        self.model_type = model_type
        self.layers: List[Tuple[torch.Tensor, torch.Tensor]] = []
        for W, b in layers:
            Wt = torch.as_tensor(np.asarray(W) if not isinstance(W, torch.Tensor) else W)
            bt = torch.as_tensor(np.asarray(b) if not isinstance(b, torch.Tensor) else b)
            self.layers.append((Wt.detach().to(self.device, torch.float32).contiguous(),
                                bt.detach().to(self.device, torch.float32).contiguous()))
        for (W0, _), (W1, _) in zip(self.layers[:-1], self.layers[1:]):
            assert W1.shape[1] == W0.shape[0], "layer shapes do not chain"
        self.input_dim = self.layers[0][0].shape[1]
        self.n_output_neurons = self.layers[-1][0].shape[0]
        # "fused": the whole predict_proba in ONE launch (lmi_mlp_topn) where the shape allows, else "tc";
        # "tc": tcgen05 tf32 split-precision GEMM per layer + softmax/top-n launch (the check of "fused", and the
        # path of every other shape); "ffma": fp32 CUDA-core kernel (verification, and the fallback for input
        # widths that are not a multiple of 4)
        self.impl = os.environ.get("LMI_MLP_IMPL", "fused")
        self._w_parts: Optional[List[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]]] = None
        self._fused_ws: dict = {}  # nq -> zero-initialised workspace of lmi_mlp_topn

    def launches_per_topn(self, x_is_f16: bool = True, n_top: int = 1) -> int:
        """Kernel launches of one ``topn_device`` call: 1 fused, else input split + one per layer + softmax/top-n."""
        if self._fused_ok(x_is_f16, n_top):
            return 1
        return (1 if self._tc_ok() else 0) + len(self.layers) + 1

    # -- construction from the reference's objects ------------------------------------
    @classmethod
    def from_reference(cls, nn, device=None) -> "NeuralNetwork":
        """From a reference ``NeuralNetwork`` (``nn.model.layers`` is the ``Sequential`` of
        ``Linear``/``ReLU``, model.py:94) or from anything exposing ``state_dict()``."""
        if isinstance(nn, cls):
            return nn
        if isinstance(nn, (list, tuple)):  # raw [(W, b), ...]
            return cls(nn, device=device)
        module = getattr(nn, "model", nn)
        sd = module.state_dict()
        idx = sorted({int(k.split(".")[-2]) for k in sd if k.endswith(".weight")})
        prefix = "layers." if any(k.startswith("layers.") for k in sd) else ""
        layers = [(sd[f"{prefix}{i}.weight"], sd[f"{prefix}{i}.bias"]) for i in idx]
        return cls(layers, device=device)

    # -- pickling (search/li/utils.py:14-29 saves the whole index with pickle) -----------
    def __getstate__(self):
        return {"layers": [(W.cpu().numpy(), b.cpu().numpy()) for W, b in self.layers], "model_type": self.model_type,
                "impl": self.impl}

    def __setstate__(self, state):
        self.__init__(state["layers"], state.get("model_type"))
        self.impl = state.get("impl", self.impl)

    # -- kernels ----------------------------------------------------------------------
    def _split(self, t: torch.Tensor, exact_in_one_part: bool):
        """hi/mid/lo tf32 parts of a device tensor (``lmi_split_tf32``)."""
        L = _native.lib()
        hi = torch.empty(t.shape, dtype=torch.float32, device=self.device)
        mid = lo = None
        if not exact_in_one_part:
            mid, lo = torch.empty_like(hi), torch.empty_like(hi)
        _native.check(L.lmi_split_tf32(t.data_ptr(), int(t.dtype == torch.float16), t.numel(), hi.data_ptr(),
                                       mid.data_ptr() if mid is not None else None,
                                       lo.data_ptr() if lo is not None else None, _stream_ptr()), "lmi_split_tf32")
        return hi, mid, lo

    def _tc_ok(self) -> bool:
        return self.impl in ("tc", "fused") and all(W.shape[1] % 4 == 0 for W, _ in self.layers)

    def _fused_ok(self, x_is_f16: bool, n_top: int) -> bool:
        if self.impl != "fused" or len(self.layers) != 2:
            return False
        (W1, _), (W2, _) = self.layers
        return bool(_native.lib().lmi_mlp_topn_supported(int(x_is_f16), W1.shape[1], W1.shape[0], W2.shape[0],
                                                         int(n_top)))

    def _weight_parts(self):
        if self._w_parts is None:
            self._w_parts = [self._split(W, False) for W, _ in self.layers]
        return self._w_parts

    def _fused(self, x: torch.Tensor, n_top: int, want_logits: bool = False):
        """One launch of ``lmi_mlp_topn``: (order, probs, logits-or-None)."""
        L = _native.lib()
        nq = x.shape[0]
        (W1, b1), (W2, b2) = self.layers
        H, C = W1.shape[0], W2.shape[0]
        ws = self._fused_ws.get(nq)
        if ws is None:
            if len(self._fused_ws) >= 16:
                self._fused_ws.pop(next(iter(self._fused_ws)))
            ws = torch.zeros(int(L.lmi_mlp_topn_workspace_bytes(nq, H, C)), dtype=torch.uint8, device=self.device)
            self._fused_ws[nq] = ws
        (h1, m1, l1), (h2, m2, l2) = self._weight_parts()
        order = probs = logits = None
        if n_top > 0:
            order = torch.empty((nq, n_top), dtype=torch.int32, device=self.device)
            probs = torch.empty((nq, n_top), dtype=torch.float32, device=self.device)
        if want_logits or n_top == 0:
            logits = torch.empty((nq, C), dtype=torch.float32, device=self.device)
        ptr = lambda t: t.data_ptr() if t is not None else None
        _native.check(L.lmi_mlp_topn(x.data_ptr(), nq, W1.shape[1], h1.data_ptr(), m1.data_ptr(), l1.data_ptr(),
                                     b1.data_ptr(), H, h2.data_ptr(), m2.data_ptr(), l2.data_ptr(), b2.data_ptr(), C,
                                     int(n_top), ws.data_ptr(), ptr(order), ptr(probs), ptr(logits), _stream_ptr()),
                      "lmi_mlp_topn")
        return order, probs, logits

    def split_input(self, x: torch.Tensor):
        """The tf32 parts of a query matrix for ``logits(x, parts=...)`` -- shared by every model of
        a multi-level index (None when this model runs the FFMA kernel)."""
        if x.shape[0] == 0 or not self._tc_ok():
            return None
        return self._split(x, x.dtype == torch.float16)  # fp16 values are exact in one tf32 part

    def logits(self, x: torch.Tensor, parts=None) -> torch.Tensor:
        """``Model.forward`` (model.py:97-99) on a device matrix (fp16 or f32)."""
        L = _native.lib()
        assert x.is_cuda and x.dim() == 2 and x.is_contiguous()
        assert x.shape[1] == self.input_dim, (x.shape, self.input_dim)
        nq = x.shape[0]
        if 0 < nq <= FUSED_LOGITS_MAX_ROWS and parts is None and self._fused_ok(x.dtype == torch.float16, 0):
            return self._fused(x, 0)[2]
        if nq > 0 and self._tc_ok():
            self._weight_parts()
            ptr = lambda t: t.data_ptr() if t is not None else None
            if parts is None:
                parts = self.split_input(x)
            out = None
            for li, ((W, b), wp) in enumerate(zip(self.layers, self._w_parts)):
                last = li + 1 == len(self.layers)
                shape = (nq, W.shape[0])
                if last:
                    out = torch.empty(shape, dtype=torch.float32, device=self.device)
                    nxt = (None, None, None)
                else:
                    nxt = tuple(torch.empty(shape, dtype=torch.float32, device=self.device) for _ in range(3))
                _native.check(L.lmi_linear_tc(ptr(parts[0]), ptr(parts[1]), ptr(parts[2]), ptr(wp[0]), ptr(wp[1]),
                                              ptr(wp[2]), b.data_ptr(), nq, W.shape[1], W.shape[0], int(not last),
                                              ptr(out) if last else None, ptr(nxt[0]), ptr(nxt[1]), ptr(nxt[2]),
                                              _stream_ptr()), "lmi_linear_tc")
                parts = nxt
            return out
        h = x
        for li, (W, b) in enumerate(self.layers):
            out = torch.empty((nq, W.shape[0]), dtype=torch.float32, device=self.device)
            _native.check(L.lmi_linear(h.data_ptr(), int(h.dtype == torch.float16), W.data_ptr(), b.data_ptr(),
                                       out.data_ptr(), nq, W.shape[1], W.shape[0],
                                       int(li + 1 < len(self.layers)), _stream_ptr()), "lmi_linear")
            h = out
        return h

    def topn_device(self, x: torch.Tensor, n_top: int) -> Tuple[torch.Tensor, torch.Tensor]:
        """(bucket ids int32[nq,n_top], probabilities f32[nq,n_top]) on the device."""
        L = _native.lib()
        assert x.is_cuda and x.dim() == 2 and x.is_contiguous() and x.shape[1] == self.input_dim
        if x.shape[0] > 0 and n_top > 0 and self._fused_ok(x.dtype == torch.float16, n_top):
            order, probs, _ = self._fused(x, n_top)
            return order, probs
        logits = self.logits(x)
        nq, C = logits.shape
        order = torch.empty((nq, n_top), dtype=torch.int32, device=self.device)
        probs = torch.empty((nq, n_top), dtype=torch.float32, device=self.device)
        _native.check(L.lmi_softmax_topn(logits.data_ptr(), nq, C, n_top, order.data_ptr(), probs.data_ptr(),
                                         _stream_ptr()), "lmi_softmax_topn")
        return order, probs

    def predict_proba(self, data_X) -> Tuple[np.ndarray, np.ndarray]:
        """search/li/model.py:226-241: all classes, most probable first."""
        x = to_device_matrix(data_X, self.device)
        order, probs = self.topn_device(x, self.n_output_neurons)
        return probs.cpu().numpy(), order.cpu().numpy().astype(np.int64)

    def predict_device(self, x: torch.Tensor, chunk: int = 1 << 20) -> torch.Tensor:
        L = _native.lib()
        out = torch.empty(x.shape[0], dtype=torch.int64, device=self.device)
        for s in range(0, x.shape[0], chunk):
            xs = x[s : s + chunk]
            logits = self.logits(xs)
            _native.check(L.lmi_argmax(logits.data_ptr(), logits.shape[0], logits.shape[1],
                                       out[s : s + chunk].data_ptr(), _stream_ptr()), "lmi_argmax")
        return out

    def predict(self, data_X) -> np.ndarray:
        """search/li/model.py:213-224: argmax class per row."""
        return self.predict_device(to_device_matrix(data_X, self.device)).cpu().numpy()

"""Host-side mirror of the DVAE codebook search directly behind the log-mel path.

Reference (file:line in /root/reference):

* ``Quantize``                       ttts/vqvae/xtts_dvae.py:52-130   (forward in eval mode :72-127, embed_code :129-130)
* ``DiscreteVAE.get_codebook_indices`` xtts_dvae.py:327-334           ``sampled, codes, _ = self.codebook(logits)``
* caller                             ttts/prepare/extract_vq.py:24-29 (mel -> codes -> ``.melvq.pth``)

Same class name, constructor arguments, buffers (``embed [dim, n_embed]``, ``cluster_size``, ``embed_avg``) and return
orders as the reference.  What runs on the GPU is the search ``argmax(-dist)`` (tcgen05 kernel ``k_vq_tc``), the
``embed_code`` gather, ``diff`` and the straight-through value, behind ``include/ttts_vq.h``.  The EMA codebook update
of training mode (:104-117) and the balancing heuristic (:73-85) belong to the trainer, not to this path: they raise.
CUDA tensors only -- there is no CPU path here.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.nn as nn

from . import _cabi


class VQPlan:
    """Device-side constants of one codebook (operand tiles split hi / lo, |e|^2, the transposed table).
    Immutable; ``nearest`` is re-entrant across streams."""

    def __init__(self, embed: torch.Tensor, device: torch.device):
        if embed.dim() != 2:
            raise ValueError("embed must be [dim, n_embed]")
        if device.type != "cuda":
            raise RuntimeError("the codebook search needs a CUDA device (sm_100a); there is no CPU fallback")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = device
        self.dim, self.n_embed = (int(s) for s in embed.shape)
        host = embed.detach().to(device="cpu", dtype=torch.float32).contiguous()
        self._lib = _cabi.lib()          # the library this plan lives in (tests may select the lab build)
        self._h = None
        h = ctypes.c_void_p()
        with torch.cuda.device(device):
            _cabi.check(self._lib.ttts_vq_plan_create(host.data_ptr(), self.dim, self.n_embed, ctypes.byref(h)))
        self._h = h

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.ttts_vq_plan_destroy(h)
            except Exception:
                pass

    def _flat(self, z: torch.Tensor) -> torch.Tensor:
        if z.device != self.device:
            raise RuntimeError(f"input on {z.device}, plan on {self.device}")
        if z.dtype != torch.float32:
            raise TypeError(f"expected float32 input, got {z.dtype}")
        flat = z.reshape(-1, self.dim)
        if flat.stride(-1) != 1 or (flat.shape[0] > 1 and flat.stride(0) < self.dim):
            flat = flat.contiguous()
        return flat

    def nearest(self, z: torch.Tensor, want_quantize: bool = False, want_diff: bool = False):
        """z ``[..., dim]`` float32 CUDA -> (embed_ind int64 ``[...]``, quantize ``[..., dim]`` or None, diff scalar
        tensor or None)."""
        flat = self._flat(z)
        n = int(flat.shape[0])
        if n == 0:
            raise RuntimeError("empty input")
        lib = self._lib
        with torch.cuda.device(self.device):
            ind = torch.empty(n, dtype=torch.int64, device=self.device)
            quant = torch.empty((n, self.dim), dtype=torch.float32, device=self.device) if want_quantize else None
            diff = torch.empty((), dtype=torch.float32, device=self.device) if want_diff else None
            ws_bytes = int(lib.ttts_vq_workspace_bytes(self._h, n))
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=self.device)
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _cabi.check(lib.ttts_vq_nearest(self._h, flat.data_ptr(), int(flat.stride(0)) if n > 1 else self.dim, n,
                                            ind.data_ptr(), quant.data_ptr() if quant is not None else None,
                                            diff.data_ptr() if diff is not None else None,
                                            ws.data_ptr(), ws_bytes, stream))
            ws.record_stream(torch.cuda.current_stream(self.device))
        shape = tuple(z.shape[:-1])
        return ind.view(shape), (quant.view(*shape, self.dim) if quant is not None else None), diff

    def nearest_bct(self, z: torch.Tensor) -> torch.Tensor:
        """z ``[B, dim, T]`` float32 CUDA (the encoder's own layout; any batch / channel strides, unit stride along T)
        -> embed_ind int64 ``[B, T]``: the search without the permute + reshape copy."""
        if z.device != self.device:
            raise RuntimeError(f"input on {z.device}, plan on {self.device}")
        if z.dtype != torch.float32:
            raise TypeError(f"expected float32 input, got {z.dtype}")
        if z.dim() != 3 or z.shape[1] != self.dim:
            raise RuntimeError(f"expected [B, {self.dim}, T], got {tuple(z.shape)}")
        B, _, T = (int(v) for v in z.shape)
        if B == 0 or T == 0:
            raise RuntimeError("empty input")
        if z.stride(2) != 1 or z.stride(1) < T or z.stride(0) < z.stride(1) * self.dim:
            z = z.contiguous()
        lib = self._lib
        with torch.cuda.device(self.device):
            ind = torch.empty((B, T), dtype=torch.int64, device=self.device)
            ws_bytes = int(lib.ttts_vq_workspace_bytes(self._h, B * T))
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=self.device)
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _cabi.check(lib.ttts_vq_nearest_bct(self._h, z.data_ptr(), int(z.stride(0)), int(z.stride(1)), B, T,
                                                ind.data_ptr(), ws.data_ptr(), ws_bytes, stream))
            ws.record_stream(torch.cuda.current_stream(self.device))
        return ind

    def embed_code(self, embed_id: torch.Tensor) -> torch.Tensor:
        if embed_id.device != self.device:
            raise RuntimeError(f"indices on {embed_id.device}, plan on {self.device}")
        ids = embed_id.to(torch.int64).reshape(-1).contiguous()
        n = int(ids.numel())
        if n == 0:
            raise RuntimeError("empty input")
        with torch.cuda.device(self.device):
            out = torch.empty((n, self.dim), dtype=torch.float32, device=self.device)
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _cabi.check(self._lib.ttts_vq_embed_code(self._h, ids.data_ptr(), n, out.data_ptr(), self.dim, stream))
        return out.view(*embed_id.shape, self.dim)

    def status(self) -> None:
        """Synchronise and raise if a kernel of this plan flagged an error (tests; optional for callers)."""
        with torch.cuda.device(self.device):
            _cabi.check(self._lib.ttts_vq_status(self._h, torch.cuda.current_stream(self.device).cuda_stream))


class Quantize(nn.Module):
    """Drop-in for ``ttts.vqvae.xtts_dvae.Quantize`` in eval mode (the ``get_codebook_indices`` path)."""

    def __init__(self, dim, n_embed, decay=0.99, eps=1e-5, balancing_heuristic=False, new_return_order=False):
        super().__init__()
        self.dim = dim
        self.n_embed = n_embed
        self.decay = decay
        self.eps = eps
        self.balancing_heuristic = balancing_heuristic
        self.codes = None
        self.max_codes = 64000
        self.codes_full = False
        self.new_return_order = new_return_order
        embed = torch.randn(dim, n_embed)
        self.register_buffer("embed", embed)
        self.register_buffer("cluster_size", torch.zeros(n_embed))
        self.register_buffer("embed_avg", embed.clone())
        self._plan: Optional[VQPlan] = None
        self._plan_key = None

    def plan(self, device: torch.device) -> VQPlan:
        """The device plan of the current ``embed`` buffer (rebuilt when the buffer is replaced or modified)."""
        key = (device, self.embed.data_ptr(), self.embed._version)
        if self._plan is None or self._plan_key != key:
            self._plan = VQPlan(self.embed, device)
            self._plan_key = key
        return self._plan

    def forward(self, input, return_soft_codes=False):
        if self.training:
            raise NotImplementedError("Quantize: only eval mode (the get_codebook_indices path) is built; the EMA "
                                      "codebook update of training (xtts_dvae.py:104-117) belongs to the trainer")
        if self.balancing_heuristic:
            raise NotImplementedError("Quantize: balancing_heuristic (xtts_dvae.py:73-85, 95-102) is trainer state")
        if return_soft_codes:
            raise NotImplementedError("Quantize: return_soft_codes materialises the [rows, n_embed] distances, "
                                      "which this path never writes to HBM")
        if input.device.type != "cuda":
            raise RuntimeError("the codebook search needs a CUDA device (sm_100a); there is no CPU fallback")
        if input.shape[-1] != self.dim:
            raise RuntimeError(f"shape '[-1, {self.dim}]' is invalid for input of size {input.numel()}")
        embed_ind, quantize, diff = self.plan(input.device).nearest(input, want_quantize=True, want_diff=True)
        if self.new_return_order:
            return quantize, embed_ind, diff
        return quantize, diff, embed_ind

    def embed_code(self, embed_id):
        if embed_id.device.type != "cuda":
            raise RuntimeError("the codebook gather needs a CUDA device (sm_100a); there is no CPU fallback")
        return self.plan(embed_id.device).embed_code(embed_id)


@torch.no_grad()
def get_codebook_indices(codebook: Quantize, logits: torch.Tensor) -> torch.Tensor:
    """The tail of ``DiscreteVAE.get_codebook_indices`` (xtts_dvae.py:332-334) for encoder output ``logits``
    ``[B, T', dim]`` (already permuted, :331): the codes alone, without the gather and ``diff``."""
    if logits.device.type != "cuda":
        raise RuntimeError("the codebook search needs a CUDA device (sm_100a); there is no CPU fallback")
    plan = codebook.plan(logits.device)
    if logits.dim() == 3 and logits.dtype == torch.float32 and logits.stride(1) == 1 and logits.stride(2) >= logits.shape[1] \
            and logits.shape[0] > 0 and logits.shape[1] > 1:
        # a permuted view of the encoder's [B, dim, T'] output (xtts_dvae.py:331): search it where it lies
        return plan.nearest_bct(logits.permute(0, 2, 1))
    return plan.nearest(logits)[0]

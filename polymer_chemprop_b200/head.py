"""FFN head + masked MSE loss on the encoder output through libwdmpnn.

``MoleculeModel.ffn`` stays the reference's ``nn.Sequential`` (chemprop/models/model.py:57-116: same modules, same
``ffn.<i>.weight / bias`` checkpoint keys); what changes is who runs it:

* ``ffn_mse_loss(encodings, ffn, targets, mask, ...)`` -> ``(preds, loss)``: the forward of model.py:174-194 plus the
  regression loss of chemprop/train/train.py:46-74, ``(MSE(preds, targets) * target_weights * data_weights *
  mask).sum() / mask.sum()``, as ONE C call (``wd_head_forward``), and its whole backward -- gradients of every
  ``Linear`` and of the encodings -- as ONE more (``wd_head_backward``) instead of ~25 torch kernels.
* ``ffn_predict(encodings, ffn)`` -> ``preds`` (chemprop/train/predict.py:10-68).

fp32 CUDA-core arithmetic, deterministic.  No CPU fallback.  Dropout layers must be inactive (p == 0 or ``eval()``);
classification / multiclass / spectra heads are not accelerated (``NotImplementedError``)."""
import ctypes as C
from typing import List, Optional, Tuple

import torch
import torch.nn as nn

from . import _capi

_ACT_OF = {nn.ReLU: "ReLU", nn.LeakyReLU: "LeakyReLU", nn.Tanh: "tanh", nn.SELU: "SELU", nn.ELU: "ELU"}


def _parse(ffn: nn.Sequential) -> Tuple[List[nn.Linear], str]:
    """The Linear layers of a reference-built FFN and its activation name; refuses anything create_ffn cannot build."""
    linears, act = [], None
    for m in ffn:
        if isinstance(m, nn.Linear):
            if m.bias is None:
                raise NotImplementedError("FFN layers without bias are not what create_ffn builds")
            linears.append(m)
        elif isinstance(m, nn.Dropout):
            if m.p > 0.0 and m.training:
                raise NotImplementedError("the fused FFN head runs with dropout off (p = 0 or model.eval())")
        elif type(m) in _ACT_OF:
            name = _ACT_OF[type(m)]
            if isinstance(m, nn.LeakyReLU) and abs(m.negative_slope - 0.1) > 1e-12:
                raise NotImplementedError("LeakyReLU slope must be 0.1 (chemprop/nn_utils.py:82)")
            if act not in (None, name):
                raise NotImplementedError("one activation type per FFN")
            act = name
        else:
            raise NotImplementedError(f"FFN module {type(m).__name__} is not accelerated (regression heads only)")
    if not 1 <= len(linears) <= _capi.HEAD_MAX_LAYERS:
        raise NotImplementedError(f"ffn_num_layers must be in [1, {_capi.HEAD_MAX_LAYERS}]")
    for a, b in zip(linears[:-1], linears[1:]):
        if a.out_features != b.in_features or (len(linears) > 2 and a.out_features != linears[0].out_features):
            raise ValueError("FFN layer widths do not chain")
    return linears, act or "ReLU"


def _head_struct(linears, act, weights, biases) -> _capi.WdHead:
    h = _capi.WdHead()
    h.n_layers, h.activation = len(linears), _capi.ACTIVATIONS[act]
    h.in_dim, h.out_dim = linears[0].in_features, linears[-1].out_features
    h.hidden_dim = linears[0].out_features if len(linears) > 1 else linears[0].out_features
    for i, (w, b) in enumerate(zip(weights, biases)):
        h.W[i], h.b[i] = w.data_ptr(), b.data_ptr()
    return h


class _HeadFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, targets, weights, mask, meta, *params):
        linears, act = meta
        dev = x.device
        n = len(linears)
        ws_ = [p.detach().contiguous() for p in params[:n]]
        bs_ = [p.detach().contiguous() for p in params[n:]]
        xc = x.detach().contiguous().float()
        head = _head_struct(linears, act, ws_, bs_)
        L, p = _capi.lib(), _capi.ptr
        G = xc.shape[0]
        with _capi.device_guard(dev):
            nbytes = L.wd_head_workspace_bytes(C.byref(head), G)
            if nbytes == 0:
                _capi.check(_capi.WD_ERR_INVALID_ARG)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            preds = torch.empty((G, head.out_dim), dtype=torch.float32, device=dev)
            loss = torch.empty((), dtype=torch.float32, device=dev)
            _capi.check(L.wd_head_forward(C.byref(head), p(xc), xc.stride(0), G, p(targets), p(weights), p(mask), p(ws), nbytes,
                                          p(preds), p(loss), _capi.current_stream(dev)))
        ctx.save_for_backward(xc, *params)
        ctx.meta, ctx.ws, ctx.nbytes, ctx.keep = meta, ws, nbytes, (ws_, bs_, targets, weights, mask)
        ctx.needs_x = x.requires_grad
        return preds, loss

    @staticmethod
    def backward(ctx, d_preds, d_loss):
        if ctx.ws is None:
            raise RuntimeError("the fused FFN head's backward ran twice on one forward (retain_graph is not supported)")
        linears, act = ctx.meta
        saved = ctx.saved_tensors                     # raises if a parameter was modified in place since forward
        xc, params = saved[0], saved[1:]
        n = len(linears)
        ws_ = [p.detach().contiguous() for p in params[:n]]
        bs_ = [p.detach().contiguous() for p in params[n:]]
        dev = xc.device
        head = _head_struct(linears, act, ws_, bs_)
        L, p = _capi.lib(), _capi.ptr
        G = xc.shape[0]
        with _capi.device_guard(dev):
            gW = [torch.empty_like(w) for w in ws_]
            gb = [torch.empty_like(b) for b in bs_]
            ldx = (head.in_dim + 3) // 4 * 4              # the library writes dX rows as float4
            dX = torch.empty((G, ldx), dtype=torch.float32, device=dev) if ctx.needs_x else None
            grads = _capi.WdHeadGrads()
            for i in range(n):
                grads.dW[i], grads.db[i] = gW[i].data_ptr(), gb[i].data_ptr()
            dl = d_loss.detach().contiguous().float() if d_loss is not None else None
            dp = d_preds.detach().contiguous().float() if d_preds is not None else None
            _capi.check(L.wd_head_backward(C.byref(head), p(xc), xc.stride(0), G, p(ctx.ws), ctx.nbytes, p(dl), p(dp), p(dX),
                                           ldx, C.byref(grads), _capi.current_stream(dev)))
        ctx.ws = None
        if dX is not None and ldx != head.in_dim:
            dX = dX[:, :head.in_dim]
        return (dX, None, None, None, None, *gW, *gb)


def ffn_mse_loss(encodings: torch.Tensor, ffn: nn.Sequential, targets: torch.Tensor, mask: torch.Tensor,
                 target_weights: Optional[torch.Tensor] = None, data_weights: Optional[torch.Tensor] = None):
    """``(preds, loss)`` of the reference's regression training step on ``encodings`` (the MPN output):
    ``preds = ffn(encodings)`` (model.py:174-194) and ``loss`` as train.py:70-74 forms it.  ``targets`` and ``mask`` are
    ``[G, num_tasks]`` (mask: bool or float), ``target_weights`` ``[G, num_tasks]`` or ``[num_tasks]``, ``data_weights``
    ``[G, 1]`` or ``[G]``, both optional (ones).  ``loss.backward()`` fills the gradients of every FFN parameter and flows
    into the encoder through ``encodings``."""
    if encodings.device.type != "cuda":
        raise RuntimeError("the fused FFN head runs on CUDA devices only (no CPU fallback)")
    linears, act = _parse(ffn)
    dev = encodings.device
    G, T = encodings.shape[0], linears[-1].out_features
    tg = targets.to(dev, torch.float32).reshape(G, T).contiguous()
    mk = mask.to(dev, torch.float32).reshape(G, T).contiguous()
    wt = None
    if target_weights is not None or data_weights is not None:
        wt = torch.ones((G, T), dtype=torch.float32, device=dev)
        if target_weights is not None:
            wt = wt * target_weights.to(dev, torch.float32).reshape(-1, T)
        if data_weights is not None:
            wt = wt * data_weights.to(dev, torch.float32).reshape(G, 1)
        wt = wt.contiguous()
    params = [m.weight for m in linears] + [m.bias for m in linears]
    return _HeadFunction.apply(encodings, tg, wt, mk, (linears, act), *params)


def ffn_predict(encodings: torch.Tensor, ffn: nn.Sequential) -> torch.Tensor:
    """``ffn(encodings)`` without a loss (inference: chemprop/train/predict.py:10-68); not differentiable."""
    if encodings.device.type != "cuda":
        raise RuntimeError("the fused FFN head runs on CUDA devices only (no CPU fallback)")
    linears, act = _parse(ffn)
    dev = encodings.device
    xc = encodings.detach().contiguous().float()
    ws_ = [m.weight.detach().contiguous() for m in linears]
    bs_ = [m.bias.detach().contiguous() for m in linears]
    head = _head_struct(linears, act, ws_, bs_)
    L, p = _capi.lib(), _capi.ptr
    with _capi.device_guard(dev):
        nbytes = L.wd_head_workspace_bytes(C.byref(head), xc.shape[0])
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        preds = torch.empty((xc.shape[0], head.out_dim), dtype=torch.float32, device=dev)
        _capi.check(L.wd_head_forward(C.byref(head), p(xc), xc.stride(0), xc.shape[0], None, None, None, p(ws), nbytes, p(preds),
                                      None, _capi.current_stream(dev)))
    return preds

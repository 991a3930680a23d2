"""Oracle: the backward replicating-portfolio fit, restated in torch-CPU (TEST INFRASTRUCTURE).

Follows ``Replicating_Portfolio.py:128-227`` and ``European Options.ipynb`` cells 12-13 under the
Keras 2.x semantics written out in SURVEY.md Appendix B.  TensorFlow is not installed in this
image and its init/shuffle RNG could not be reproduced anyway, so **fit parity is unpinned at bit
level by the reference**: this restatement is the checker, the reference's saved notebook numbers
are statistical anchors only.  Forward/backward go through ``torch.autograd`` (fp32), which keeps
this checker independent of the hand-derived gradients in the CUDA kernels.

Parameter vector layout shared with the CUDA path (Keras kernel shape is ``(in, out)``):
``W1[in][8], b1[8], W2[8][8], b2[8], W3[8][out], b3[out]`` with ``in, out = (1, 1)`` for the
``'eo'`` net (97 parameters, psi = 1 - phi) and ``(3, 2)`` for the ``'rp'`` net (122 parameters).
"""
from __future__ import annotations

import numpy as np
import torch

LEAKY_ALPHA = 0.3           # Keras LeakyReLU default (activation='LeakyReLU' string)
HIDDEN = 8
NET_IO = {"eo": (1, 1), "rp": (3, 2)}


def n_params(net: str) -> int:
    i, o = NET_IO[net]
    return i * HIDDEN + HIDDEN + HIDDEN * HIDDEN + HIDDEN + HIDDEN * o + o


def init_params(net: str, seed: int, out_bias) -> np.ndarray:
    """Kernels ~ N(0, 0.1^2) (``Replicating_Portfolio.py:149``), hidden biases 0, output bias
    constant (``:150``: ``[1-p_otm, p_otm]``; EO cell 12: ``payoff.mean()/S0``).  The generator is
    ours (NumPy PCG64), the reference's TF generator is not reproducible."""
    i, o = NET_IO[net]
    rng = np.random.default_rng(seed)
    w1 = rng.normal(0.0, 0.1, size=(i, HIDDEN))
    w2 = rng.normal(0.0, 0.1, size=(HIDDEN, HIDDEN))
    w3 = rng.normal(0.0, 0.1, size=(HIDDEN, o))
    b3 = np.broadcast_to(np.asarray(out_bias, dtype=np.float64), (o,))
    return np.concatenate([w1.ravel(), np.zeros(HIDDEN), w2.ravel(), np.zeros(HIDDEN),
                           w3.ravel(), b3]).astype(np.float32)


def unpack(params: torch.Tensor, net: str):
    i, o = NET_IO[net]
    ofs = 0

    def take(n, shape):
        nonlocal ofs
        t = params[ofs:ofs + n].reshape(shape)
        ofs += n
        return t

    return (take(i * HIDDEN, (i, HIDDEN)), take(HIDDEN, (HIDDEN,)),
            take(HIDDEN * HIDDEN, (HIDDEN, HIDDEN)), take(HIDDEN, (HIDDEN,)),
            take(HIDDEN * o, (HIDDEN, o)), take(o, (o,)))


def forward(params: torch.Tensor, net: str, x: torch.Tensor, prices: torch.Tensor):
    """``x``: (n, in) features at time t; ``prices``: (n, 2) stock and bond price.
    Returns ``(V, phi, psi)``."""
    w1, b1, w2, b2, w3, b3 = unpack(params, net)
    h = torch.nn.functional.leaky_relu(x @ w1 + b1, LEAKY_ALPHA)
    h = torch.nn.functional.leaky_relu(h @ w2 + b2, LEAKY_ALPHA)
    out = h @ w3 + b3
    if net == "eo":
        phi = out[:, 0]
        psi = 1.0 - phi
    else:
        phi, psi = out[:, 0], out[:, 1]
    return phi * prices[:, 0] + psi * prices[:, 1], phi, psi


def loss_fn(kind: str, y: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    if kind == "mse":
        return ((y - v) ** 2).mean()
    if kind == "pinball":                     # Replicating_Portfolio.py:138-142
        e = y - v
        return torch.maximum(0.99 * e, (0.99 - 1.0) * e).mean()
    raise ValueError(kind)


def lr_schedule(epoch: int, lr: float) -> float:
    """``Replicating_Portfolio.py:128-136``."""
    if epoch < 100:
        return 1e-2
    if epoch < 200:
        return 1e-3
    if epoch < 400:
        return 5e-4
    return lr


class AdamState:
    """Keras Adam: ``lr_t = lr sqrt(1-b2^t)/(1-b1^t)``; ``w -= lr_t m / (sqrt(v) + eps)``,
    ``eps = 1e-7``.  One instance per compiled Keras model (``:175-180``)."""

    def __init__(self, n: int, lr: float = 1e-3):
        self.m = torch.zeros(n, dtype=torch.float32)
        self.v = torch.zeros(n, dtype=torch.float32)
        self.t = 0
        self.lr = lr
        self.b1, self.b2, self.eps = 0.9, 0.999, 1e-7

    def step(self, params: torch.Tensor, grad: torch.Tensor) -> None:
        self.t += 1
        self.m.mul_(self.b1).add_(grad, alpha=1 - self.b1)
        self.v.mul_(self.b2).addcmul_(grad, grad, value=1 - self.b2)
        lr_t = self.lr * np.sqrt(1.0 - self.b2 ** self.t) / (1.0 - self.b1 ** self.t)
        params.sub_(np.float32(lr_t) * self.m / (self.v.sqrt() + np.float32(self.eps)))


def fit(params: torch.Tensor, opt: AdamState, net: str, x, prices, y, *, epochs: int,
        patience: int, use_schedule: bool, loss: str = "mse", batch_size=512,
        shuffle_rng=None, perm_fn=None):
    """One Keras ``model.fit`` call (``Replicating_Portfolio.py:211``), in place on ``params``.

    ``batch_size=None`` is one full-batch Adam step per epoch, no shuffle; the scaled mode of the CUDA path
    (default: 8 index-order steps per epoch) is ``batch_size=n // 8`` with ``shuffle_rng=None``.  Returns ``dict(epochs_run, losses, stopped_early)``.
    EarlyStopping(monitor='loss', min_delta=0, restore_best_weights=True) as in TF 2.x: the wait
    counter is bumped first and reset on improvement; the best weights come back only when the
    stop fired before the last epoch."""
    n = x.shape[0]
    best, best_w, wait = np.inf, None, 0
    losses, stopped = [], False
    for epoch in range(epochs):
        if use_schedule:
            opt.lr = lr_schedule(epoch, opt.lr)
        if batch_size is None:
            batches = [slice(None)]
        else:
            if perm_fn is not None:
                perm = torch.from_numpy(perm_fn(epoch))
            else:
                perm = torch.from_numpy(shuffle_rng.permutation(n)) if shuffle_rng is not None else torch.arange(n)
            batches = [perm[s:s + batch_size] for s in range(0, n, batch_size)]
        tot = 0.0
        for idx in batches:
            p = params.detach().clone().requires_grad_(True)
            v, _, _ = forward(p, net, x[idx], prices[idx])
            l = loss_fn(loss, y[idx], v)
            (g,) = torch.autograd.grad(l, p)
            opt.step(params, g)
            nb = n if batch_size is None else len(idx)
            tot += float(l.detach()) * nb
        cur = tot / n
        losses.append(cur)
        wait += 1
        if cur < best:
            best, wait = cur, 0
            best_w = params.clone()
        if wait >= patience and epoch > 0:
            stopped = True
            params.copy_(best_w)
            break
    return {"epochs_run": len(losses), "losses": losses, "stopped_early": stopped}


M64 = (1 << 64) - 1


def perm_pow2(x: int, bits: int, seed: int) -> int:
    """The epoch permutation of the CUDA mini-batch kernel (``perm_pow2`` in ``csrc/orp_fit.cu``):
    odd multiply, xorshift, odd multiply on ``bits`` bits.  Ours, not Keras' (its shuffle is not
    reproducible); restated here so both sides visit samples in the same order."""
    mask = (1 << bits) - 1
    a1 = ((seed * 0x9E3779B97F4A7C15) & M64) | 1
    a2 = (((seed ^ 0xD1B54A32D192ED03) * 0xBF58476D1CE4E5B9) & M64) | 1
    x = ((x * a1 + (seed >> 7)) & M64) & mask
    x ^= x >> ((bits + 1) // 2)
    x = ((x * a2) & M64) & mask
    x ^= x >> ((bits + 2) // 3)
    return x & mask


def minibatch_perm(n: int, seed: int) -> np.ndarray:
    """Sample order of one epoch: position -> sample index (cycle-walking below ``n``)."""
    bits = 0
    while (1 << bits) < n:
        bits += 1
    out = np.empty(n, dtype=np.int64)
    for pos in range(n):
        i = perm_pow2(pos, bits, seed)
        while i >= n:
            i = perm_pow2(i, bits, seed)
        out[pos] = i
    return out


def backward_replication(S_cols: np.ndarray, B_cols: np.ndarray, terminal_value: np.ndarray, *,
                         net: str = "eo", scale: float = 1.0, init_seed: int = 1234,
                         out_bias=None, batch_size=512, shuffle_seed: int = 0,
                         epochs_first: int = 500, epochs_next: int = 100,
                         patience_first: int = 50, patience_next: int = 7,
                         features=None, cuda_perm: bool = False, batch_order=None):
    """The loop ``for t_i in range(n-2, -1, -1)`` of ``European Options.ipynb`` cell 13 (model1
    only, ``values[:, t_i] = g_t``); with ``net='rp'`` the MSE half of
    ``Replicating_Portfolio.py:193-221``.

    ``S_cols``: (paths, n) kept columns; ``B_cols``: (n,); ``terminal_value``: (paths,) already in
    network units (EO: ``payoff / S0``).  ``scale`` divides prices (EO: ``ADJUSTMENT_FACTOR = S0``).
    ``features(t_i)`` supplies the extra RP inputs ``(N_t/N, lambda_t)``; default 1 and 0.
    ``batch_order`` (optional, position -> sample index, the same every epoch): the make-up of the optimiser steps
    when it is not "consecutive index ranges" — the CUDA path's virtual shards take step ``k`` from the ``k``-th range
    of every shard (``orp_b200.fit.ReplicationNet.split``).
    Returns ``dict(V0, phi0, psi0, values, history)`` in network units (times ``scale`` for EO's
    ``V0_RP``)."""
    n_paths, n = S_cols.shape
    if out_bias is None:
        out_bias = float(terminal_value.mean())
    params = torch.from_numpy(init_params(net, init_seed, out_bias).copy())
    opt = AdamState(params.numel())
    rng = np.random.default_rng(shuffle_seed) if shuffle_seed is not None else None     # None: index order, no shuffle
    values = np.empty((n_paths, n))
    values[:, -1] = terminal_value
    hist, phi0, psi0 = [], None, None
    first = True
    for t_i in range(n - 2, -1, -1):
        s_t = torch.from_numpy((S_cols[:, t_i] / scale).astype(np.float32))
        s_t1 = torch.from_numpy((S_cols[:, t_i + 1] / scale).astype(np.float32))
        b_t = torch.full_like(s_t, np.float32(B_cols[t_i] / scale))
        b_t1 = torch.full_like(s_t, np.float32(B_cols[t_i + 1] / scale))
        if net == "eo":
            x = s_t[:, None]
        else:
            extra = features(t_i) if features is not None else (np.ones(n_paths), np.zeros(n_paths))
            x = torch.stack([s_t] + [torch.from_numpy(np.asarray(e, dtype=np.float32)) for e in extra], dim=1)
        p0 = torch.stack([s_t, b_t], dim=1)
        p1 = torch.stack([s_t1, b_t1], dim=1)
        y = torch.from_numpy(values[:, t_i + 1].astype(np.float32))
        h = fit(params, opt, net, x, p1, y, epochs=epochs_first if first else epochs_next,
                patience=patience_first if first else patience_next, use_schedule=first,
                batch_size=batch_size, shuffle_rng=rng,
                perm_fn=(lambda e: np.asarray(batch_order, dtype=np.int64)) if batch_order is not None else
                (lambda e, t_i=t_i: minibatch_perm(n_paths, (shuffle_seed * 4096 + 2 * t_i) * 1_000_003 + e))
                if (cuda_perm and batch_size is not None) else None)
        with torch.no_grad():
            g_t, phi, psi = forward(params, net, x, p0)
        values[:, t_i] = g_t.numpy().astype(np.float64)
        hist.append(dict(t_i=t_i, **{k: h[k] for k in ("epochs_run", "stopped_early")}, loss=h["losses"][-1]))
        phi0, psi0 = float(phi.double().mean()), float(psi.double().mean())
        first = False
    return {"V0": float(values[:, 0].mean()), "phi0": phi0, "psi0": psi0, "values": values,
            "history": hist, "params": params.numpy().copy()}

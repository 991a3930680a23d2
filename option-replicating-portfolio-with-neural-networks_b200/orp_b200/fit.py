"""Backward per-time-step replicating-portfolio fit on the GPU (host driver).

Mirrors the loop ``for t_i in trange(n_time_steps-2, -1, -1)`` of ``Replicating_Portfolio.py:193-227`` and
``European Options.ipynb`` cell 13 with the Keras semantics of SURVEY.md Appendix B.  Two modes:

* ``mode='scaled'`` (B200-native): ``steps_per_epoch`` (default 8, the notebook's own 4096 / 512) optimiser steps per
  pass over the paths, each on 1/8 of ALL paths of all ranks — the fused forward/backward kernel fills every SM, the
  ``[gradient, loss, count]`` vectors of the ranks are exchanged inside the persistent kernel (or summed by a collective
  in the two-stage form), Adam / LR schedule / early stopping run on the device: one launch per ``Model.fit``.
* ``mode='keras'``: batch_size-512 mini-batches in a shuffled order, serial by definition, one thread block.

Weights, both Adam states and the learning rate persist across time steps exactly as the reference's
compiled models do (``:175-180, 203-211``).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
import os

from .dist import allgather_rows, allgather_tree_sum_, allreduce_sum_, is_distributed, rank_world, tree_sum_rows

HIDDEN = 8
NET_IO = {"eo": (1, 1), "rp": (3, 2)}


def n_params(net: str) -> int:
    return _lib.lib().orp_fit_n_params(_lib.NET[net])


def init_params(net: str, seed: int, out_bias) -> np.ndarray:
    """``RandomNormal(0, 0.1)`` kernels, zero hidden biases, constant output bias
    (``Replicating_Portfolio.py:149-150``; EO cell 12).  TensorFlow's generator cannot be reproduced, so the
    draw is ours: NumPy PCG64 seeded with ``seed``.  Layout: W1[in][8], b1, W2[8][8], b2, W3[8][out], b3."""
    i, o = NET_IO[net]
    rng = np.random.default_rng(seed)
    w1 = rng.normal(0.0, 0.1, size=(i, HIDDEN))
    w2 = rng.normal(0.0, 0.1, size=(HIDDEN, HIDDEN))
    w3 = rng.normal(0.0, 0.1, size=(HIDDEN, o))
    b3 = np.broadcast_to(np.asarray(out_bias, dtype=np.float64), (o,))
    return np.concatenate([w1.ravel(), np.zeros(HIDDEN), w2.ravel(), np.zeros(HIDDEN), w3.ravel(), b3]).astype(np.float32)


def scale_column(col, inv_scale: float, out=None):
    """``(float)((double) col * inv_scale)`` on the device: a price column in network units (EO cell 13 divides by
    ``ADJUSTMENT_FACTOR``; Keras casts to fp32).  The fit kernels read such fp32 columns as they are."""
    torch = _lib.require_cuda()
    if not col.is_contiguous():
        raise ValueError("columns must be contiguous (use the time-major path tensor)")
    if out is None:
        out = torch.empty(col.shape, dtype=torch.float32, device=col.device)
    with torch.cuda.device(col.device):
        _lib.check(_lib.lib().orp_fit_scale_column(_lib.ptr(col), _lib.F64 if col.dtype == torch.float64 else _lib.F32,
                                                   col.numel(), float(inv_scale), _lib.ptr(out), _lib.stream_ptr()))
    return out


def _nccl_group(group) -> bool:
    """True when torch.distributed runs over NCCL with more than one rank (one process per GPU on one box)."""
    if not is_distributed(group):
        return False
    import torch.distributed as dist
    return str(dist.get_backend(group)).lower() == "nccl"


class Optimizer:
    """Device-resident Adam + callback state of one compiled Keras model (``keras.optimizers.Adam(1e-3)``)."""

    def __init__(self, device):
        torch = _lib.require_cuda()
        self.state = torch.empty(_lib.FIT_STATE_DOUBLES, dtype=torch.float64, device=device)
        self._head_idx = None
        with torch.cuda.device(device):
            _lib.check(_lib.lib().orp_fit_init(_lib.ptr(self.state), _lib.stream_ptr()))

    def head(self) -> np.ndarray:
        """The 16 scalar control words plus (last entry) the barrier-abort flag, in ONE device-to-host copy."""
        if self._head_idx is None:
            import torch
            self._head_idx = torch.tensor(list(range(16)) + [_lib.ST_ABORT], device=self.state.device)
        return self.state[self._head_idx].cpu().numpy()


_peer_cache = {}          # (device, process group) -> mapped exchange buffers: created once per process, not per run


class ReplicationNet:
    """The shared weights of model1/model2 plus scratch buffers; ``fit`` runs one ``Model.fit`` call."""

    def __init__(self, net: str, params0: np.ndarray, device="cuda", group=None, peer_exchange=None):
        torch = _lib.require_cuda()
        if net not in _lib.NET:
            raise ValueError(f"net must be one of {sorted(_lib.NET)}")
        self.net, self.device, self.group = net, torch.device(device), group
        self.P = n_params(net)
        if params0.shape != (self.P,):
            raise ValueError(f"expected {self.P} parameters for net '{net}'")
        self.params = torch.zeros(_lib.FIT_MAX_PARAMS, dtype=torch.float32, device=self.device)
        self.params[:self.P] = torch.from_numpy(np.ascontiguousarray(params0, dtype=np.float32)).to(self.device)
        self.gsum = torch.zeros(_lib.GSUM_LEN, dtype=torch.float64, device=self.device)
        # zero-filled: the persistent kernel's hand-offs are tagged with a step counter kept in this buffer
        self.scratch = torch.zeros(_lib.lib().orp_fit_scratch_bytes(), dtype=torch.uint8, device=self.device)
        self.sums = torch.zeros((1, 5), dtype=torch.float64, device=self.device)
        self.launches = 0
        self.two_stage = False      # True forces grad -> [all-reduce] -> update even on one GPU (tests, multi-GPU form)
        self.persistent = True      # False: one launch per optimiser step (orp_fit_epochs) instead of one per fit
        self.peers = None           # multi-GPU: exchange buffers mapped over CUDA IPC (see _setup_peers)
        self._equal_counts = {}
        if peer_exchange is None:
            peer_exchange = _nccl_group(group) and os.environ.get("ORP_B200_NO_PEERS") is None
        if peer_exchange:
            # the exchange buffers (a cudaMalloc, the IPC handle exchange and the mappings) are made once per process,
            # device and group and shared by every model created afterwards: their step counter lives in the buffer
            key = (str(self.device), id(group) if group is not None else None)
            if key not in _peer_cache:
                self._setup_peers()
                _peer_cache[key] = self.peers
            self.peers = _peer_cache[key]

    def _setup_peers(self):
        """Exchange buffers for the in-kernel all-gather of the [gradient, loss, count] vectors: every rank creates
        one (the library cudaMallocs it: CUDA IPC needs its own base allocation), the 64-byte IPC handles travel
        through torch.distributed, and each rank maps the others' buffers.  Single process: a world of one.

        The exchange needs every rank on ONE host with peer access between all GPUs.  The ranks agree on that before
        anything is mapped (host names and ``can_device_access_peer`` are gathered, the outcome of the mapping is
        all-reduced); if any rank cannot take part, every rank undoes its mappings and the fit falls back to the
        two-stage form (``orp_fit_grad`` -> collective -> ``orp_fit_update``), which needs nothing of the kind."""
        torch = _lib.require_cuda()
        L = _lib.lib()
        rank, world = rank_world(self.group)
        with torch.cuda.device(self.device):
            if world > 1:
                import socket
                import torch.distributed as dist
                dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
                info = [None] * world
                dist.all_gather_object(info, (socket.gethostname(), dev_index), group=self.group)
                same_host = all(h == info[0][0] for h, _ in info)
                reach = same_host and all(d == dev_index or torch.cuda.can_device_access_peer(dev_index, d) for _, d in info)
                flag = torch.tensor([1.0 if reach else 0.0], device=self.device)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
                if float(flag) == 0.0:
                    self.peers = None
                    return
            own = C.c_void_p()
            handle = (C.c_ubyte * 64)()
            _lib.check(L.orp_peer_buffer_create(C.byref(own), handle))
            ptrs, opened, ok = [None] * world, [], True
            if world > 1:
                handles = [None] * world
                dist.all_gather_object(handles, bytes(handle), group=self.group)
                for r, h in enumerate(handles):
                    if r == rank:
                        continue
                    q = C.c_void_p()
                    buf = (C.c_ubyte * 64).from_buffer_copy(h)
                    if L.orp_peer_buffer_open(buf, C.byref(q)) != _lib.ORP_OK:
                        ok = False
                        break
                    ptrs[r] = q.value
                    opened.append(q.value)
                flag = torch.tensor([1.0 if ok else 0.0], device=self.device)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)     # also: nobody starts before every mapping exists
                if float(flag) == 0.0:
                    for q in opened:
                        L.orp_peer_buffer_close(q)
                    dist.barrier(group=self.group)                                # every rank has unmapped before anyone frees
                    L.orp_peer_buffer_destroy(own.value)
                    self.peers = None
                    return
            ptrs[rank] = own.value
            table = torch.tensor(ptrs, dtype=torch.int64, device=self.device)
        self.peers = dict(rank=rank, world=world, own=own.value, opened=opened, table=table)

    def close(self):
        """Detach from the exchange buffers (they stay mapped for the next model of this process; release_peers() frees
        them)."""
        self.peers = None

    def _ranks_have_equal_counts(self, count: int) -> bool:
        if count not in self._equal_counts:
            torch = _lib.require_cuda()
            if self.peers["world"] == 1:
                self._equal_counts[count] = True
            else:
                import torch.distributed as dist
                t = torch.tensor([count, -count], dtype=torch.int64, device=self.device)
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
                self._equal_counts[count] = int(t[0]) == -int(t[1])
        return self._equal_counts[count]

    def batch(self, s_t, s_t1, target, inv_scale, bond_t, bond_t1, loss="mse", feat1=None, feat2=None,
              skip_nonfinite: bool = False, segments: int = 1):
        """One time step's inputs as the C struct.  fp32 price columns that are not in network units yet
        (``inv_scale != 1``) are converted once here (``scale_column``), because the kernels take fp32 columns as they
        are.  ``segments``: the arrays are cut into that many equally long pieces that are reduced separately and
        added pairwise (``orp_fit_batch.n_seg``: the virtual shards that make the sums rank-count invariant)."""
        torch = _lib.require_cuda()
        for t in (s_t, s_t1, target, feat1, feat2):
            if t is not None and not t.is_contiguous():
                raise ValueError("columns must be contiguous (use the time-major path tensor)")
        if s_t.dtype != torch.float64 and float(inv_scale) != 1.0:
            s_t, s_t1 = scale_column(s_t, inv_scale), scale_column(s_t1, inv_scale)
            bond_t, bond_t1, inv_scale = bond_t * float(inv_scale), bond_t1 * float(inv_scale), 1.0
        n, L = s_t.numel(), int(segments)
        if L < 1 or L > _lib.FIT_MAX_SEGMENTS or n % L:
            raise ValueError(f"segments must be in [1, {_lib.FIT_MAX_SEGMENTS}] and divide the path count")
        b = _lib.FitBatch()
        b.net, b.loss = _lib.NET[self.net], _lib.LOSS[loss]
        b.paths_dtype = _lib.F64 if s_t.dtype == torch.float64 else _lib.F32
        b.reserved = 1 if skip_nonfinite else 0          # ORP_FIT_SKIP_NONFINITE
        b.count = n // L
        b.n_seg, b.seg_stride = L, n // L
        b.d_s_t, b.d_s_t1 = _lib.ptr(s_t), _lib.ptr(s_t1)
        b.d_feat1, b.d_feat2, b.d_target = _lib.ptr(feat1) or None, _lib.ptr(feat2) or None, _lib.ptr(target) or None
        b.inv_scale, b.bond_t, b.bond_t1 = float(inv_scale), float(bond_t), float(bond_t1)
        b._keep = (s_t, s_t1, target, feat1, feat2)          # keep the tensors alive
        b._esz = s_t.element_size()
        return b

    @staticmethod
    def _offset(b, off: int, count: int, n_seg: int):
        """A copy of ``b`` whose arrays start ``off`` samples later: ``n_seg`` segments of ``count`` samples."""
        c = _lib.FitBatch()
        C.memmove(C.byref(c), C.byref(b), C.sizeof(_lib.FitBatch))
        c.count, c.n_seg = count, n_seg
        c.d_s_t, c.d_s_t1 = b.d_s_t + off * b._esz, b.d_s_t1 + off * b._esz
        c.d_target = b.d_target + off * 4 if b.d_target else None
        c.d_feat1 = b.d_feat1 + off * 4 if b.d_feat1 else None
        c.d_feat2 = b.d_feat2 + off * 4 if b.d_feat2 else None
        c._keep, c._esz = b._keep, b._esz
        return c

    @staticmethod
    def split(b, steps_per_epoch: int):
        """The batch cut into ``steps_per_epoch`` optimiser steps (Keras' ``batch_size`` = ``count / steps_per_epoch``):
        step ``k`` takes the ``k``-th contiguous range of EVERY segment.  Contiguous blocks of a Sobol sequence are
        themselves well stratified, so no shuffle is needed; on several GPUs step ``k`` is the union of the ``k``-th
        ranges of all virtual shards, whichever rank holds them."""
        n = int(b.count)
        k = max(1, min(int(steps_per_epoch), n))
        if k == 1:
            return [b]
        per = -(-n // k)
        return [ReplicationNet._offset(b, off, min(per, n - off), int(b.n_seg) or 1) for off in range(0, n, per)]

    def fit(self, batch, opt: Optimizer, epochs: int, patience: int, use_schedule: bool, mode: str = "scaled",
            batch_size: int = 512, shuffle_seed: int = 0, check_every: int = 16, steps_per_epoch: int = 1):
        """One ``Model.fit`` (``Replicating_Portfolio.py:211``).  ``steps_per_epoch`` (scaled mode): optimiser
        steps per pass over the paths.  Returns ``dict(epochs_run, loss, stopped_early, lr)``."""
        torch = _lib.require_cuda()
        L = _lib.lib()
        netc = _lib.NET[self.net]
        with torch.cuda.device(self.device):
            st = _lib.stream_ptr()
            _lib.check(L.orp_fit_begin(_lib.ptr(opt.state), int(epochs), int(patience), int(bool(use_schedule)), st))
            parts = self.split(batch, steps_per_epoch) if mode == "scaled" else [batch]
            dist_on = is_distributed(self.group)
            done = 0
            single_launch = mode == "scaled" and not dist_on and not self.two_stage
            if (mode == "scaled" and self.peers is not None and not self.two_stage
                    and self._ranks_have_equal_counts(int(batch.count))):
                # one process per GPU: the same persistent kernel; the ranks' sums are exchanged inside it over NVLink
                _lib.check(L.orp_fit_run_peers(C.byref(batch), int(steps_per_epoch), _lib.ptr(self.params), _lib.ptr(opt.state),
                                               _lib.ptr(self.gsum), _lib.ptr(self.scratch), int(epochs),
                                               self.peers["rank"], self.peers["world"], _lib.ptr(self.peers["table"]), st))
                self.launches += 1
                done = epochs
            elif single_launch and self.persistent:
                # one process, one GPU: the whole epoch loop is ONE cooperative launch (tagged hand-overs between blocks);
                # the kernel leaves as soon as EarlyStopping fires, so there is nothing to poll
                _lib.check(L.orp_fit_run(C.byref(batch), int(steps_per_epoch), _lib.ptr(self.params), _lib.ptr(opt.state),
                                         _lib.ptr(self.gsum), _lib.ptr(self.scratch), int(epochs), st))
                self.launches += 1
                done = epochs
            elif single_launch:
                parts_arr = (_lib.FitBatch * len(parts))()
                for k, part in enumerate(parts):
                    C.memmove(C.byref(parts_arr[k]), C.byref(part), C.sizeof(_lib.FitBatch))
            while done < epochs:
                n = min(check_every, epochs - done)
                if single_launch:
                    # one process, one GPU: grad + fixed-order reduction + Adam/callbacks in ONE launch per optimiser
                    # step, and the epoch loop itself runs inside the library (no Python between launches)
                    _lib.check(L.orp_fit_epochs(parts_arr, len(parts), _lib.ptr(self.params), _lib.ptr(opt.state),
                                                _lib.ptr(self.gsum), _lib.ptr(self.scratch), n, st))
                    self.launches += n * len(parts)
                    done += n
                    if opt.state[_lib.ST_STOP].item() != 0.0:
                        break
                    continue
                for e in range(n):
                    if mode == "scaled":
                        for k, part in enumerate(parts):
                            _lib.check(L.orp_fit_grad(C.byref(part), _lib.ptr(self.params), _lib.ptr(opt.state),
                                                      _lib.ptr(self.gsum), _lib.ptr(self.scratch), st))
                            if dist_on:
                                allgather_tree_sum_(self.gsum, self.group)     # pairwise in rank order, as the peer exchange adds
                            flags = (1 if k == 0 else 0) | (2 if k == len(parts) - 1 else 0)
                            _lib.check(L.orp_fit_update(netc, _lib.ptr(self.params), _lib.ptr(opt.state),
                                                        _lib.ptr(self.gsum), flags, st))
                        self.launches += 3 * len(parts)
                    elif mode == "keras":
                        if is_distributed(self.group):
                            raise ValueError("mode='keras' is serial by definition: replicas only, no sharding")
                        _lib.check(L.orp_fit_epoch_minibatch(C.byref(batch), _lib.ptr(self.params), _lib.ptr(opt.state),
                                                             int(batch_size), int(shuffle_seed) * 1_000_003 + done + e, st))
                        self.launches += 1
                    else:
                        raise ValueError("mode must be 'scaled' or 'keras'")
                done += n
                if opt.state[_lib.ST_STOP].item() != 0.0:       # one 8-byte read per `check_every` epochs
                    break
            _lib.check(L.orp_fit_end(netc, _lib.ptr(self.params), _lib.ptr(opt.state), st))
            self.launches += 2
            h = opt.head()
            if h[-1] != 0.0:
                raise RuntimeError("orp_fit_run: a hand-over between thread blocks (or ranks) timed out; "
                                   "the optimiser state is unusable — create a new Optimizer")
        return {"epochs_run": int(h[_lib.ST_EPOCHS_RUN]), "loss": float(h[_lib.ST_LAST_LOSS]),
                "stopped_early": bool(h[_lib.ST_STOPPED_EARLY]), "lr": float(h[_lib.ST_LR])}

    def residuals(self, batch, out):
        """``out[i] = target_i - V(X1)_i`` with the current weights: the reference's per-path ``VaR`` array
        (``get_phi_psi_VaR``, ``Replicating_Portfolio.py:120``), kept on the device."""
        torch = _lib.require_cuda()
        L, n = int(batch.n_seg) or 1, int(batch.count)
        with torch.cuda.device(self.device):
            for v in range(L):
                seg = self._offset(batch, v * int(batch.seg_stride), n, 1) if L > 1 else batch
                _lib.check(_lib.lib().orp_fit_residuals(C.byref(seg), _lib.ptr(self.params), _lib.ptr(out) + 4 * v * int(batch.seg_stride),
                                                        _lib.stream_ptr()))
                self.launches += 1
        return out

    def predict(self, batch, values_out=None):
        """``model.predict(X0)`` into ``values_out`` (fp32) and the global sums
        ``(sum V, sum phi, sum psi, sum |y - V(X1)|, sum |y - V(X1)| / max(|y|, 1e-7))`` as a 5-vector on the host.
        One kernel pass per segment; the segments' sums of all ranks are added pairwise in global segment order on the
        host, so the result does not depend on the number of ranks that hold the virtual shards."""
        torch = _lib.require_cuda()
        L, n = int(batch.n_seg) or 1, int(batch.count)
        if self.sums.shape[0] != L:
            self.sums = torch.zeros((L, 5), dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            for v in range(L):
                seg = self._offset(batch, v * int(batch.seg_stride), n, 1) if L > 1 else batch
                vo = _lib.ptr(values_out) + 4 * v * int(batch.seg_stride) if values_out is not None else 0
                _lib.check(_lib.lib().orp_fit_predict(C.byref(seg), _lib.ptr(self.params), vo, _lib.ptr(self.sums[v]),
                                                      _lib.ptr(self.scratch), _lib.stream_ptr()))
                self.launches += 2
            rows = allgather_rows(self.sums, self.group)             # (world, L, 5): global segment v = rank * L + local
            return tree_sum_rows(rows.reshape(-1, 5))


def release_peers():
    """Unmap the peers' exchange buffers and free this rank's (collective: call on every rank, e.g. before
    ``destroy_process_group``).  Optional: process exit releases them as well."""
    torch = _lib.require_cuda()
    L = _lib.lib()
    for (dev, _), peers in list(_peer_cache.items()):
        if peers is None:
            continue
        with torch.cuda.device(torch.device(dev)):
            torch.cuda.synchronize()
            if peers["world"] > 1:
                import torch.distributed as dist
                dist.barrier()                        # no rank frees memory a peer's kernel may still write to
            for q in peers["opened"]:
                L.orp_peer_buffer_close(q)
            L.orp_peer_buffer_destroy(peers["own"])
    _peer_cache.clear()


def virtual_segments(count: int, group=None, virtual_shards: int = 8) -> int:
    """How many of the ``virtual_shards`` equal pieces of the GLOBAL path range this rank holds: ``virtual_shards /
    world`` when that is a whole number and divides the local path count (world = 1, 2, 4, 8 for the default 8), else
    1 (the sums are then still deterministic, but depend on the world size)."""
    _, world = rank_world(group)
    if virtual_shards >= world and virtual_shards % world == 0:
        L = virtual_shards // world
        if L <= _lib.FIT_MAX_SEGMENTS and count % L == 0 and count >= L:
            return L
    return 1


def fit_replication(paths_tm, bond, terminal_values, *, net: str = "eo", mode: str = "scaled", scale: float = 1.0,
                    out_bias=None, init_seed: int = 1234, total_paths: int | None = None, group=None,
                    epochs_first: int = 500, epochs_next: int = 100, patience_first: int = 50, patience_next: int = 7,
                    batch_size: int = 512, shuffle_seed: int = 0, features=None, second_model: bool = False,
                    cost_of_capital: float = 0.1, check_every: int = 16, verbose: bool = False, on_step=None,
                    steps_per_epoch: int = 8, var_quantiles=None, on_step_end=None, skip_nonfinite: bool = False,
                    virtual_shards: int | None = None):
    """Backward replication over all kept columns.

    ``paths_tm``: TIME-MAJOR ``(n_cols, count)`` CUDA tensor of this rank's paths; ``bond``: ``n_cols`` bond
    prices; ``terminal_values``: fp32 ``(count,)`` CUDA tensor in network units (EO: ``payoff / S0``).
    ``scale`` = ``ADJUSTMENT_FACTOR`` dividing prices (EO cell 13: ``S0``).  ``features(t_i)`` returns the two
    extra fp32 inputs of the ``'rp'`` net ``(N_t/N, lambda_t)`` or ``None``.
    ``steps_per_epoch`` (scaled mode): Adam steps per pass over the paths; the default 8 is the reference
    notebook's own ratio (4096 paths / batch 512, ``European Options.ipynb`` cell 13), kept at every scale.  ``second_model=True`` adds the
    pinball-loss fit of the SAME weights and the blend ``values = g + cost_of_capital (h - g)``
    (``Replicating_Portfolio.py:217-221``; the reference's model2 aliases model1's layers).

    ``var_quantiles`` (e.g. ``(.98, .99)``): per step, the quantiles of the replication residual
    ``values[:, t+1] - phi Y_{t+1} - psi B_{t+1}`` over ALL paths of all ranks (``Replicating_Portfolio.py:120-122``),
    by radix select on the device; reported as ``steps[i]["VaR"]``.  ``on_step_end(rec)`` is called with each record.

    ``skip_nonfinite`` (opt-in, not the reference's behaviour): paths whose features, prices or target are NaN / inf
    take no part in the fits, keep NaN as their value and stay out of the means and quantiles — without it one such
    path turns every weight into NaN, exactly as in the reference.

    ``virtual_shards`` (scaled mode): the global path range is treated as that many equal pieces; every optimiser step,
    predict sum and mean is formed per piece and the pieces are added pairwise in index order, so ``V0``, the hedge
    weights and the fitted parameters are bit-identical on 1, 2, 4 or 8 GPUs (``virtual_segments``).  Default: 8 from
    2^18 paths (all ranks together) on, 1 below.  The price of the invariance is the make-up of an optimiser step: step
    ``k`` is the ``k``-th range of every shard instead of one contiguous index range, i.e. 8 aligned Sobol blocks of 1/64
    of the paths instead of one block of 1/8 — immaterial for large sets (2^20 paths: 8 x 16384-point nets per step),
    but at the reference's 4096 paths it would replace a 512-point net by 8 clustered 64-point nets and visibly raise
    the gradient noise, so small runs keep the contiguous batches.

    Returns ``dict(V0, phi0, psi0, steps=[...], launches)`` in network units."""
    torch = _lib.require_cuda()
    n_cols, count = paths_tm.shape
    device = paths_tm.device
    bond = np.asarray(bond, dtype=np.float64)
    if bond.shape != (n_cols,):
        raise ValueError("bond must have one entry per kept column")
    if virtual_shards is None:
        if total_paths is None:
            cnt = torch.tensor([float(count)], dtype=torch.float64, device=device)
            n_global = float(allreduce_sum_(cnt, group))
        else:
            n_global = float(total_paths)
        virtual_shards = 8 if n_global >= 2 ** 18 else 1
    L = virtual_segments(count, group, int(virtual_shards)) if mode == "scaled" else 1
    seg = count // L
    tv = terminal_values.contiguous()
    rows = torch.zeros((L, 2), dtype=torch.float64, device=device)      # per virtual shard: (paths that count, sum of targets)
    for v in range(L):
        piece = tv[v * seg:(v + 1) * seg].double()
        if skip_nonfinite:
            fin = torch.isfinite(piece) & torch.isfinite(paths_tm[-1, v * seg:(v + 1) * seg])     # e.g. -inf price, payoff 0
            rows[v, 0], rows[v, 1] = fin.sum(), piece[fin].sum()
        else:
            rows[v, 0], rows[v, 1] = float(seg), piece.sum()
    tot = tree_sum_rows(allgather_rows(rows, group).reshape(-1, 2))    # pairwise over the global shards: rank-count invariant
    total = float(tot[0]) if total_paths is None else float(total_paths)
    if out_bias is None:
        out_bias = tot[1] / total                      # EO cell 12: payoff.mean() / ADJUSTMENT_FACTOR
    model = ReplicationNet(net, init_params(net, init_seed, out_bias), device, group)
    opt1 = Optimizer(device)
    opt2 = Optimizer(device) if second_model else None
    v_next = tv
    v_cur = torch.empty_like(v_next)
    h_buf = torch.empty_like(v_next) if second_model else None
    steps, first = [], True
    resid_buf = None
    phi0 = psi0 = v0 = float("nan")
    # Price columns in network units, fp32, converted ONCE per time step (column t+1 of this step is column t of the
    # previous one): the epoch loop then reads 12 bytes per sample and does no fp64 arithmetic.  Same bits as the
    # kernels' own conversion of the fp64 columns.
    inv = 1.0 / scale
    x_next = scale_column(paths_tm[n_cols - 1], inv)
    x_cur = torch.empty_like(x_next)
    for t_i in range(n_cols - 2, -1, -1):
        if on_step is not None:
            on_step({"t_i": t_i})
        f1, f2 = features(t_i) if features is not None else (None, None)
        scale_column(paths_tm[t_i], inv, out=x_cur)
        common = dict(inv_scale=1.0, bond_t=bond[t_i] * inv, bond_t1=bond[t_i + 1] * inv, feat1=f1, feat2=f2,
                      skip_nonfinite=skip_nonfinite, segments=L)
        model.launches += 1
        b = model.batch(x_cur, x_next, v_next, loss="mse", **common)
        h = model.fit(b, opt1, epochs_first if first else epochs_next, patience_first if first else patience_next,
                      use_schedule=first, mode=mode, batch_size=batch_size,
                      shuffle_seed=shuffle_seed * 4096 + 2 * t_i, check_every=check_every, steps_per_epoch=steps_per_epoch)
        sums = model.predict(b, v_cur)
        if skip_nonfinite:                              # the means are over the paths that took part (NaN marks the others)
            n_ok = torch.isfinite(v_cur).sum().double().reshape(1)
            allreduce_sum_(n_ok, group)
            total = max(float(n_ok), 1.0)
        rec = dict(t_i=t_i, epochs_run=h["epochs_run"], loss=h["loss"], stopped_early=h["stopped_early"],
                   mean_V=sums[0] / total, phi=sums[1] / total, psi=sums[2] / total, mae=sums[3] / total,
                   mape=100.0 * sums[4] / total)       # Keras 'mae' / 'mape' of model1.evaluate(X1, ...), RP.py:215
        if second_model:
            b2 = model.batch(x_cur, x_next, v_next, loss="pinball", **common)
            h2 = model.fit(b2, opt2, epochs_first if first else epochs_next, patience_first if first else patience_next,
                           use_schedule=first, mode=mode, batch_size=batch_size,
                           shuffle_seed=shuffle_seed * 4096 + 2 * t_i + 1, check_every=check_every,
                           steps_per_epoch=steps_per_epoch)
            sums2 = model.predict(b2, h_buf)
            v_cur.add_(h_buf - v_cur, alpha=cost_of_capital)     # values = g + cc (h - g)
            # get_phi_psi_VaR reads the shared 'Phi_Psi' layer after BOTH fits (Replicating_Portfolio.py:104-115)
            rec.update(epochs_run2=h2["epochs_run"], loss2=h2["loss"], phi=sums2[1] / total, psi=sums2[2] / total,
                       mean_V=sums[0] / total + cost_of_capital * (sums2[0] - sums[0]) / total)
        if var_quantiles is not None:
            from .risk import quantiles
            if resid_buf is None:
                resid_buf = torch.empty_like(v_next)
            model.residuals(b2 if second_model else b, resid_buf)          # weights as they stand after the step's fits
            rec["VaR"] = quantiles(resid_buf, var_quantiles, group, ignore_nan=skip_nonfinite, total=int(total)).tolist()
        steps.append(rec)
        if on_step_end is not None:
            on_step_end(rec)
        if verbose:
            print(rec)
        v_next, v_cur = v_cur, v_next
        x_next, x_cur = x_cur, x_next
        phi0, psi0, v0 = rec["phi"], rec["psi"], rec["mean_V"]
        first = False
    out = {"V0": v0, "phi0": phi0, "psi0": psi0, "steps": steps, "launches": model.launches,
           "params": model.params[:model.P].cpu().numpy(), "values0": v_next,
           "virtual_shards": int(virtual_shards) if mode == "scaled" else 1}
    model.close()
    return out

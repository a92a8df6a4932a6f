"""Keyframe-sharded data-parallel mapping: host-side plumbing only.

Every rank holds the full (replicated) Gaussian map and Adam state, renders a DIFFERENT keyframe of
the selected window per iteration, the per-parameter gradients live in ONE flat fp32 buffer, a single
all-reduce (NCCL over NVLink/NVSwitch on the GPU box, gloo in the CPU tests) averages it, and every
rank applies the identical Adam step -- replicas stay bit-identical without any broadcast.  The
reference has no distributed code; its sequential analogue is one uniformly drawn keyframe
per iteration (python_ordinary/scripts/splatam.py:1415-1426).
"""
import ctypes as C

import torch
import torch.distributed as dist


def keyframe_for(iteration, rank, world, window):
    """Keyframe slot rendered by `rank` at `iteration`: `world` different draws from the window per
    iteration, round-robin so that every rank sees every keyframe."""
    return (iteration * world + rank) % window


def allreduce_mean_(flat, group=None):
    """In-place average of the flat gradient buffer over all ranks (sum all-reduce, then 1/world).
    Returns the work-free tensor; a single collective covers every parameter group."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return flat
    if dist.get_backend(group) == "nccl":
        dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group)   # ncclAvg: the 1/world scaling rides in the collective
    else:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.mul_(1.0 / world)
    return flat


def replicas_identical(tensors, group=None):
    """True on every rank iff the given tensors are bit-identical across ranks (a debugging / test aid:
    compares a 64-bit checksum of the raw bits with max and min all-reduces)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return True
    sums = []
    for t in tensors:
        b = t.detach().contiguous().view(torch.int32).to(torch.int64)
        w = torch.arange(1, b.numel() + 1, dtype=torch.int64, device=b.device)
        sums.append(((b.reshape(-1) * w) % 2147483629).sum())
    s = torch.stack(sums)
    hi, lo = s.clone(), s.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    return bool(torch.equal(hi, lo))


def slice_bounds(numel, world, rank):
    """[begin, end) in floats of the slice `rank` reduces in the two-shot peer all-reduce (csrc/peer.cu):
    float4 granularity, ceil(n4 / world) float4 per rank, the tail ranks may own less or nothing."""
    n4 = numel // 4
    per = (n4 + world - 1) // world
    b = min(n4, per * rank)
    return 4 * b, 4 * min(n4, b + per)


class _RawCudaBuffer:
    """Lets torch view cudaMalloc'd memory that this package owns (CUDA array interface, no copy)."""

    def __init__(self, ptr, numel):
        self.__cuda_array_interface__ = dict(shape=(int(numel),), typestr="<f4", data=(int(ptr), False), version=2)


class PeerAllReduce:
    """The gradient average of keyframe-sharded mapping as ONE kernel over NVLink peer memory
    (isogs_allreduce_avg_p2p + isogs_peer_wait, csrc/peer.cu) instead of an NCCL all-reduce.

        ar = PeerAllReduce(device)                       # after dist.init_process_group
        gmap = GaussianMap(params, device, grad_alloc=ar.alloc)
        map_step(..., allreduce=ar)

    `alloc(numel)` is COLLECTIVE: every rank cudaMallocs its flat gradient buffer (+ 256 B of flags), the IPC
    handles travel through torch.distributed, and every rank maps all peers.  Calling the object averages the
    given prefix of the buffer in place on the current stream; every replica receives identical bits (the owner
    of a slice sums in rank order and stores the same values everywhere)."""
    FLAG_BYTES = 256

    def __init__(self, device, group=None, n_ctas=0):
        from . import _lib
        self._lib = _lib
        self.group = group
        self.device = torch.device(device)
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.n_ctas = int(n_ctas)
        self.epoch = 0
        self.status = torch.zeros((1,), dtype=torch.int32, device=self.device)
        self.own = None
        self.ptrs = []
        self.tensor = None

    def _last_error(self):
        msg = self._lib.lib().isogs_last_error()
        return msg.decode() if msg else ""

    def _agree(self, ok):
        """True iff `ok` on every rank (so that a failure on one rank makes all of them fall back together)."""
        if self.world == 1:
            return bool(ok)
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return bool(t.item())

    def _shared(self, nbytes):
        """COLLECTIVE: cudaMalloc `nbytes` (zeroed) on every rank, exchange the IPC handles, map every peer.
        Returns (own pointer, [pointer of every rank's allocation, own entry = own pointer])."""
        L = self._lib.lib()
        ptr, handle = C.c_void_p(), (C.c_ubyte * 64)()
        with torch.cuda.device(self.device):
            rc = L.isogs_peer_alloc(int(nbytes), C.byref(ptr), handle)
            err = None if rc == 0 else self._last_error()
            if not self._agree(rc == 0):
                if rc == 0:
                    L.isogs_peer_free(ptr)
                raise RuntimeError(f"isogs_peer_alloc failed on some rank ({err})")
            own = int(ptr.value)
            handles = [None] * self.world
            if self.world > 1:
                dist.all_gather_object(handles, bytes(handle), group=self.group)
            ptrs, ok, err = [], True, None
            for r in range(self.world):
                if r == self.rank:
                    ptrs.append(own)
                    continue
                q = C.c_void_p()
                hb = (C.c_ubyte * 64).from_buffer_copy(handles[r])
                if L.isogs_peer_open(hb, C.byref(q)) != 0:
                    ok, err = False, self._last_error()
                    ptrs.append(0)
                else:
                    ptrs.append(int(q.value))
            if not self._agree(ok):
                self._unshare(own, ptrs)
                raise RuntimeError(f"isogs_peer_open failed on some rank ({err})")
        return own, ptrs

    def _unshare(self, own, ptrs):
        L = self._lib.lib()
        for r, p in enumerate(ptrs):
            if r != self.rank and p:
                L.isogs_peer_close(C.c_void_p(p))
        L.isogs_peer_free(C.c_void_p(own))

    def alloc(self, numel):
        self.release()
        numel = (int(numel) + 3) & ~3
        flag_off = (4 * numel + 255) & ~255
        self.own, self.ptrs = self._shared(flag_off + self.FLAG_BYTES)
        self.bufs = (C.c_void_p * self.world)(*self.ptrs)
        self.flags = (C.c_void_p * self.world)(*[p + flag_off for p in self.ptrs])
        self.numel = numel
        self.epoch = 0
        self.tensor = torch.as_tensor(_RawCudaBuffer(self.own, numel), device=self.device)
        return self.tensor

    def __call__(self, flat):
        L = self._lib.lib()
        n = flat.numel()
        if flat.data_ptr() != self.own or n > self.numel or n % 4:
            raise ValueError("PeerAllReduce: pass a prefix (a multiple of 4 floats) of the buffer alloc() returned")
        self.epoch += 1
        st = torch.cuda.current_stream(self.device).cuda_stream
        self._lib.check(L.isogs_allreduce_avg_p2p(self.world, self.rank, self.bufs, self.flags, n, self.epoch, self.n_ctas,
                                                  self.status.data_ptr(), st))
        self._lib.check(L.isogs_peer_wait(self.world, self.flags[self.rank], self.epoch, self.status.data_ptr(), st))
        return flat

    def check_status(self):
        """One host read: raises if a peer spin timed out since the last check."""
        s = int(self.status.item())
        if s:
            self.status.zero_()
            raise RuntimeError(f"peer all-reduce timed out waiting for a rank (stage {s})")

    def release(self):
        if self.own is None:
            return
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)           # nobody may still be reading this rank's buffer
        self.tensor = None
        self._unshare(self.own, self.ptrs)
        self.own, self.ptrs = None, []


class PeerReduceAdam(PeerAllReduce):
    """The exchange FUSED with the optimizer step (isogs_reduce_adam_p2p, csrc/peer.cu): the owner of a 1/world slice
    averages it over the ranks, applies Adam to it in registers and stores the updated PARAMETERS into every rank's
    parameter tensors.  The parameter tensors of the map therefore live in peer-mapped memory too:

        ar = PeerReduceAdam(device)
        gmap = GaussianMap(params, device, grad_alloc=ar.alloc)
        ar.adopt(gmap)                                   # COLLECTIVE: moves gmap.p[...] into shared allocations
        map_step(..., allreduce=ar)                      # backward, then this object: no isogs_adam_fused launch

    Optimizer state is SHARDED: a rank keeps Adam moments only for the slice it owns.  reset_optimizer() (every mapping
    call of the reference, splatam.py:1406) is slice-agnostic; anything that moves rows (prune / densify compaction) or
    changes the Gaussian count in the middle of a mapping call must be preceded by sync_moments(gmap)."""

    def __init__(self, device, group=None, n_ctas=0):
        super().__init__(device, group, n_ctas)
        self.pown, self.pptrs = {}, {}

    def adopt(self, gmap):
        """COLLECTIVE.  Re-homes every parameter tensor that has a gradient into IPC-shared memory (same values)."""
        self.drop_params()
        for k in gmap.grad_names:
            old = gmap.p[k]
            n = old.numel()
            own, ptrs = self._shared(4 * n + 64)         # slack: the tail float4 of a group may run 3 floats over
            t = torch.as_tensor(_RawCudaBuffer(own, n), device=self.device).view(old.shape)
            t.copy_(old)
            gmap.p[k] = t
            self.pown[k], self.pptrs[k] = own, ptrs
        gmap._cs = None
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)
        return gmap

    def fused_adam(self, gmap, lrs, eps=1e-15, beta1=0.9, beta2=0.999):
        L = self._lib.lib()
        n = gmap.grad_used
        if gmap.grad_flat.data_ptr() != self.own or n > self.numel or n % 4:
            raise ValueError("PeerReduceAdam: the map's gradient buffer must come from alloc()")
        names = gmap.grad_names
        arr = (self._lib.PeerAdamGroup * len(names))()
        for i, k in enumerate(names):
            if k not in self.pptrs or gmap.p[k].data_ptr() != self.pown[k]:
                raise ValueError("PeerReduceAdam: call adopt(gmap) first (and again after the pool has grown)")
            g = gmap.g[k]
            arr[i].offset = (g.data_ptr() - gmap.grad_flat.data_ptr()) // 4
            arr[i].numel = g.numel()
            for r in range(self.world):
                arr[i].param[r] = self.pptrs[k][r]
            arr[i].exp_avg, arr[i].exp_avg_sq = gmap.m[k].data_ptr(), gmap.v[k].data_ptr()
            arr[i].lr = lrs[k]
        self.epoch += 1
        st = torch.cuda.current_stream(self.device).cuda_stream
        self._launch(L, n, arr, names, beta1, beta2, eps, gmap, st)
        self._lib.check(L.isogs_peer_wait(self.world, self.flags[self.rank], self.epoch, self.status.data_ptr(), st))

    def _launch(self, L, n, arr, names, beta1, beta2, eps, gmap, st):
        self._lib.check(L.isogs_reduce_adam_p2p(self.world, self.rank, self.bufs, self.flags, n, arr, len(names), beta1, beta2,
                                                eps, gmap.adam_state.data_ptr(), self.epoch, self.n_ctas,
                                                self.status.data_ptr(), st))

    def sync_moments(self, gmap):
        """COLLECTIVE.  Makes every rank's Adam moments complete again (each rank broadcasts the parts it owns):
        needed before rows move (prune / densify) or the ownership boundaries shift (count change) mid-call."""
        if self.world == 1:
            return
        n = gmap.grad_used
        base = gmap.grad_flat.data_ptr()
        for r in range(self.world):
            b, e = slice_bounds(n, self.world, r)
            for k in gmap.grad_names:
                off = (gmap.g[k].data_ptr() - base) // 4
                lo, hi = max(b, off) - off, min(e, off + gmap.g[k].numel()) - off
                if hi <= lo:
                    continue
                for t in (gmap.m[k], gmap.v[k]):
                    dist.broadcast(t.view(-1)[lo:hi], src=r, group=self.group)

    def drop_params(self):
        if not self.pown:
            return
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)
        for k in list(self.pown):
            self._unshare(self.pown[k], self.pptrs[k])
        self.pown, self.pptrs = {}, {}


class NvlsReduceAdam(PeerReduceAdam):
    """PeerReduceAdam through the NVSwitch's in-fabric reduction (isogs_reduce_adam_nvls): the gradient buffer and the
    parameter tensors are SYMMETRIC allocations bound to a multicast object, so the owner of a slice reads it with one
    multimem.ld_reduce (the switch adds the ranks' copies) and broadcasts the updated parameters with one multimem.st.
    The allocation / multicast binding / handle exchange is torch.distributed._symmetric_memory (plumbing: device memory
    and rendezvous); the kernel is ours.  Raises at alloc() -- on every rank together -- when the box has no multicast
    support, so the caller can fall back to PeerReduceAdam."""

    def __init__(self, device, group=None, n_ctas=0):
        super().__init__(device, group, n_ctas)
        self._symm = {}          # own pointer -> (tensor, handle, multicast pointer)

    def _shared(self, nbytes):
        import torch.distributed._symmetric_memory as symm_mem
        grp = self.group if self.group is not None else dist.group.WORLD
        ok, err, t, hdl, mc = True, None, None, None, 0
        try:
            try:
                symm_mem.set_backend("CUDA")
            except Exception:
                pass
            with torch.cuda.device(self.device):
                t = symm_mem.empty((int(nbytes) + 3) // 4, dtype=torch.float32, device=self.device)
                t.zero_()
                hdl = symm_mem.rendezvous(t, grp)
            mc = int(hdl.multicast_ptr)
            off = int(getattr(hdl, "offset", 0))
            ptrs = [int(p) + off for p in hdl.buffer_ptrs]
            if not mc:
                ok, err = False, "no multicast support (multicast_ptr == 0)"
            elif ptrs[self.rank] != t.data_ptr():
                ok, err = False, f"symmetric buffer pointer mismatch ({ptrs[self.rank]:#x} vs {t.data_ptr():#x}, offset {off})"
            mc += off
        except Exception as e:                               # noqa: BLE001 -- any failure means "no NVLS here"
            ok, err = False, f"{type(e).__name__}: {e}"
        if not self._agree(ok):
            raise RuntimeError(f"NVLS symmetric allocation unavailable on some rank ({err})")
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)                       # every rank's zero fill has landed before anybody's flags are used
        self._symm[t.data_ptr()] = (t, hdl, mc)
        return t.data_ptr(), ptrs

    def _unshare(self, own, ptrs):
        self._symm.pop(own, None)                            # the symmetric allocation is freed with its tensor

    def _launch(self, L, n, arr, names, beta1, beta2, eps, gmap, st):
        mcp = (C.c_void_p * len(names))(*[self._symm[self.pown[k]][2] for k in names])
        self._lib.check(L.isogs_reduce_adam_nvls(self.world, self.rank, self.bufs, self.flags, n, arr, len(names), beta1,
                                                 beta2, eps, gmap.adam_state.data_ptr(), self.epoch, self.n_ctas,
                                                 self.status.data_ptr(), self._symm[self.own][2], mcp, st))

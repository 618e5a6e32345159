"""View-parallel plumbing: one process per GPU, full Gaussian replica on every
rank, each rank rasterizes its own cameras, one all-reduce of the flat parameter gradient per step.

The reference has no distributed code at all; this is the only multi-GPU shape its hot path admits
(a single image is never split).  torch.distributed (NCCL on GPUs, gloo in the CPU tests) does the
transport; nothing here touches the rasterizer.
"""
import torch
import torch.distributed as dist

# parameter order and floats per Gaussian of the flat buffer: xyz 3, SH 48 (f_dc 3 + f_rest 45),
# opacity 1, scaling 3, rotation 4 = 59  (layout of the optimizer groups, scene/gaussian_model_ht.py:275-287)
PARAM_ORDER = ("means3D", "shs", "opacities", "scales", "rotations")
FLOATS_PER_GAUSSIAN = 59


def shard_views(num_views, rank, world):
    """Contiguous, balanced split of view indices: rank r gets views [lo, hi)."""
    base, extra = divmod(num_views, world)
    lo = rank * base + min(rank, extra)
    return range(lo, lo + base + (1 if rank < extra else 0))


class FlatGradReducer:
    """Packs the five parameter gradients into one contiguous buffer, all-reduces it once (sum),
    and hands the reduced views back -- one collective of 236 B/Gaussian instead of five."""

    def __init__(self, P, device, dtype=torch.float32, peer_memory=False):
        """peer_memory=True: the flat buffer lives in torch symmetric memory (mapped by every rank of the default
        process group) and `all_reduce` runs this library's own two-shot kernel over NVLink peer loads/stores
        (htgs_allreduce_p2p) between two cross-GPU barriers instead of calling NCCL.  CUDA + initialised process group
        required; construction raises if the peers cannot map each other's memory."""
        self.P = P
        n = P * FLOATS_PER_GAUSSIAN
        self._peer = None
        self.collective = "nccl"
        if peer_memory:
            import ctypes
            import torch.distributed._symmetric_memory as symm_mem
            if dtype != torch.float32:
                raise RuntimeError("peer-memory all-reduce is float32 only")
            n_pad = (n + 3) // 4 * 4  # the kernel moves 128-bit words
            buf = symm_mem.empty(n_pad, dtype=torch.float32, device=device)
            hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
            buf.zero_()
            off = buf.data_ptr() - int(hdl.buffer_ptrs[hdl.rank])
            table = (ctypes.c_void_p * hdl.world_size)(*[int(p) + off for p in hdl.buffer_ptrs])
            # a handle without NVSwitch multicast reports multicast_ptr == 0; that address is the only capability gate
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            self._peer = dict(buf=buf, hdl=hdl, table=table, n_pad=n_pad, mc=(mc + off) if mc else 0)
            # peer loads move 2(N-1)/N buffers per link direction, the switch multicast about one: NVLS from 3 ranks up
            self.collective = "multimem" if (mc and hdl.world_size > 2 and peer_memory != "p2p") or peer_memory == "multimem" else "p2p"
            if self.collective == "multimem" and not mc:
                raise RuntimeError("multimem all-reduce requested but the symmetric-memory handle has no multicast address")
            self.flat = buf[:n]
        else:
            self.flat = torch.zeros(n, device=device, dtype=dtype)
        self.slices = {}
        o = 0
        for name, per in zip(PARAM_ORDER, (3, 48, 1, 3, 4)):
            self.slices[name] = (o, o + P * per)
            o += P * per
        assert o == self.flat.numel()

    def install_arena(self, params=None):
        """Let the rasterizer backward write its gradients directly into this buffer (no pack copy).

        The backward OVERWRITES the slice, so it is only right when the slice is not still some parameter's live .grad:
        one backward per parameter per step, and `.grad = None` / `zero_grad(set_to_none=True)` between steps.  Pass
        `params` ({name in PARAM_ORDER: leaf tensor handed to the rasterizer}) to have that checked on every backward:
        a parameter that still holds its arena slice gets an ordinary accumulate instead (correct, one extra add)."""
        from diff_gaussian_rasterization import _C
        _C.set_grad_arena(self.flat, self.slices, params)

    @staticmethod
    def remove_arena():
        from diff_gaussian_rasterization import _C
        _C.set_grad_arena(None, None)

    def pack_if_needed(self, grads):
        """Copies only the gradients that autograd did not adopt in place from the arena; returns how many were copied."""
        copied = 0
        for name in PARAM_ORDER:
            lo, hi = self.slices[name]
            g = grads[name]
            if g is None:
                self.flat[lo:hi].zero_()
                copied += 1
            elif g.data_ptr() != self.flat.data_ptr() + 4 * lo:
                self.flat[lo:hi].copy_(g.reshape(-1))
                copied += 1
        return copied

    def pack(self, grads):
        for name in PARAM_ORDER:
            lo, hi = self.slices[name]
            g = grads[name]
            if g is None:
                self.flat[lo:hi].zero_()
            else:
                self.flat[lo:hi].copy_(g.reshape(-1))
        return self.flat

    def all_reduce(self, average=False):
        if self._peer is not None:
            import ctypes
            import htgs_native as nat
            pm = self._peer
            dev = self.flat.device
            with torch.cuda.device(dev):
                st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
                pm["hdl"].barrier(channel=0)  # every rank's gradients are complete (stream-ordered, device-side)
                if self.collective == "multimem":
                    nat.check(nat.lib().htgs_allreduce_multimem(ctypes.c_void_p(pm["mc"]), pm["hdl"].rank, pm["hdl"].world_size,
                                                                pm["n_pad"], st), "htgs_allreduce_multimem")
                else:
                    nat.check(nat.lib().htgs_allreduce_p2p(pm["table"], pm["hdl"].rank, pm["hdl"].world_size, pm["n_pad"], st),
                              "htgs_allreduce_p2p")
                pm["hdl"].barrier(channel=0)  # every rank's slice has been written everywhere
            if average:
                self.flat.div_(dist.get_world_size())
            return self.flat
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
            if average:
                self.flat.div_(dist.get_world_size())
        return self.flat

    def unpack(self, like):
        return {name: self.flat[lo:hi].view_as(like[name]) for name, (lo, hi) in self.slices.items()}


def reduce_densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom):
    """Densification statistics for one data-parallel step ("things that must also be
    reduced"): every rank contributes the view it rendered, every replica ends with the same state a single
    process would hold after visiting all those views one after another
    (ht3dgs_trainer.py:143-148 + gaussian_model_ht.py:718-721 per view):

        xyz_gradient_accum += sum over ranks of ||viewspace_grad[:, :2]|| where radii > 0      (SUM)
        denom              += number of ranks that saw the Gaussian                            (SUM)
        max_radii2D         = max(max_radii2D, max over ranks of radii where radii > 0)        (MAX)

    Per-rank increments are built first (on CUDA by the fused statistics kernel into zeroed scratch), then one
    SUM all-reduce of 8 B/Gaussian and one MAX all-reduce of 4 B/Gaussian, then applied in place.  Without an
    initialised process group (or world size 1) it degenerates to the single-view update."""
    P = int(radii.numel())
    dev = radii.device
    inc = torch.zeros(2, P, device=dev, dtype=torch.float32)  # [0] gradient-norm increment, [1] visit count
    rad = torch.zeros(P, device=dev, dtype=torch.float32)
    if radii.is_cuda:
        from htgs_optim import add_densification_stats
        add_densification_stats(radii.contiguous(), viewspace_grad.contiguous(), rad, inc[0], inc[1])
    else:  # host tensors: the gloo tests of this logic
        vis = radii > 0
        inc[0] = torch.where(vis, torch.linalg.vector_norm(viewspace_grad[:, :2], dim=-1), inc[0])
        inc[1] = vis.to(torch.float32)
        rad = torch.where(vis, radii.to(torch.float32), rad)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(inc, op=dist.ReduceOp.SUM)
        dist.all_reduce(rad, op=dist.ReduceOp.MAX)
    xyz_gradient_accum.add_(inc[0].view_as(xyz_gradient_accum))
    denom.add_(inc[1].view_as(denom))
    torch.maximum(max_radii2D, rad.view_as(max_radii2D), out=max_radii2D)


# ---- compact SH exchange: ranks trade the colour gradient, not the expanded SH gradient ----
# 11 floats/Gaussian are all-reduced; the 48 SH floats are rebuilt from 3 floats per view.  Rotations first: the kernels
# write dL_drotations with 128-bit stores, so its slice must start 16-byte aligned for every P.
SMALL_ORDER = ("rotations", "means3D", "opacities", "scales")
SMALL_FLOATS = (4, 3, 1, 3)


class CompactSHGradReducer:
    """View-parallel gradient exchange that never ships dL_dsh (81 % of the flat 59-float gradient).

    Per view, dL_dsh[i] is the outer product of the SH basis of Gaussian i's view direction with the clamp-masked colour
    gradient r[i] (3 floats).  Each rank's backward writes r and the camera position of its views into view SLOTS
    (diff_gaussian_rasterization/_C.py compact mode -> htgs_backward_compact); `all_reduce()` then
      1. all-reduces the remaining 11 floats/Gaussian (xyz, opacity, scaling, rotation), and
      2. rebuilds sum_views basis (x) r on every rank from ALL views' slots, in view order (htgs_sh_expand), so replicas
         stay bit-identical,
    moving 44 + 12 * views bytes per Gaussian instead of 236.  Transports for the slots:
      peer_memory="multimem"  the backward pushes its slot through the NVSwitch multicast mapping into every rank's buffer;
      peer_memory="p2p"       slots stay local, the rebuild reads the other ranks' slots with NVLink peer loads;
      peer_memory=False       torch.distributed all_gather of the slots (NCCL; any backend).
    `views_total` slots exist per step; rank r owns views [r * views_total / world, (r + 1) * views_total / world).
    Usage per step: begin_step(); forward/backward of each local view (parameters' .grad for xyz/opacity/scaling/rotation
    land in the small flat buffer through the gradient arena); all_reduce(D); grads(like) returns all five gradients."""

    def __init__(self, P, device, views_total=None, peer_memory=False, M=16):
        import ctypes
        import htgs_native as nat
        self.P, self.M, self.device = int(P), int(M), device
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        self.views_total = int(views_total) if views_total is not None else self.world
        if self.views_total % self.world:
            raise RuntimeError(f"{self.views_total} views do not divide over {self.world} ranks")
        if self.views_total > 16:
            raise RuntimeError("at most 16 view slots per step")
        self.views_per_rank = self.views_total // self.world
        self.slot_floats = int(nat.lib().htgs_sh_slot_floats(self.P))
        n_small = 11 * self.P
        self.n_small_pad = (n_small + 3) // 4 * 4
        total = self.n_small_pad + self.views_total * self.slot_floats
        self._peer = None
        self.collective = "nccl"
        if peer_memory and self.world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            buf = symm_mem.empty(total, dtype=torch.float32, device=device)
            hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
            buf.zero_()
            off = buf.data_ptr() - int(hdl.buffer_ptrs[hdl.rank])
            ptrs = [int(p) + off for p in hdl.buffer_ptrs]
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            want_mc = peer_memory == "multimem" or (peer_memory is True and mc and hdl.world_size > 2)
            if want_mc and not mc:
                raise RuntimeError("multimem transport requested but the symmetric-memory handle has no multicast address")
            self.collective = "multimem" if want_mc else "p2p"
            self._peer = dict(buf=buf, hdl=hdl, ptrs=ptrs, table=(ctypes.c_void_p * hdl.world_size)(*ptrs), mc=(mc + off) if mc else 0)
            self.buf = buf
        else:
            self.buf = torch.zeros(total, device=device, dtype=torch.float32)
        self.flat = self.buf[:n_small]
        self.slices = {}
        o = 0
        for name, per in zip(SMALL_ORDER, SMALL_FLOATS):
            self.slices[name] = (o, o + self.P * per)
            o += self.P * per
        self.sh_grad = torch.zeros((self.P, self.M, 3), device=device, dtype=torch.float32)
        self._next_local = 0
        self._side = None  # second stream of the exchange (created on first use)

    # -- slots --
    def _slot_offset(self, view):
        return self.n_small_pad + view * self.slot_floats

    def slot(self, view):
        o = self._slot_offset(view)
        return self.buf[o:o + self.slot_floats]

    def begin_step(self):
        self._next_local = 0

    def _provider(self, dev):
        if self._next_local >= self.views_per_rank:
            raise RuntimeError("more backward passes in this step than view slots owned by this rank; call begin_step()")
        view = self.rank * self.views_per_rank + self._next_local
        self._next_local += 1
        off = 4 * self._slot_offset(view)
        if self.collective == "multimem":
            return self._peer["mc"] + off, True
        return self.buf.data_ptr() + off, False

    def install(self, params=None):
        """Routes the rasterizer backward into this reducer: small gradients into the flat buffer (arena), colour
        gradients into the view slots (compact SH mode)."""
        from diff_gaussian_rasterization import _C
        owners = {k: v for k, v in params.items() if k in SMALL_ORDER} if params else None
        _C.set_grad_arena(self.flat, self.slices, owners)
        _C.set_sh_compact(self._provider)

    @staticmethod
    def remove():
        from diff_gaussian_rasterization import _C
        _C.set_grad_arena(None, None)
        _C.set_sh_compact(None)

    def pack_if_needed(self, grads):
        copied = 0
        for name in SMALL_ORDER:
            lo, hi = self.slices[name]
            g = grads[name]
            if g is None:
                self.flat[lo:hi].zero_()
                copied += 1
            elif g.data_ptr() != self.flat.data_ptr() + 4 * lo:
                self.flat[lo:hi].copy_(g.reshape(-1))
                copied += 1
        return copied

    def all_reduce(self, means3D, sh_degree):
        """After every local view's backward: leaves the summed small gradients in `flat` and the summed dL_dsh in
        `sh_grad` on every rank.  means3D: the [P,3] positions the views were rendered with."""
        if self._next_local != self.views_per_rank:
            raise RuntimeError(f"{self._next_local} of {self.views_per_rank} local views ran a backward this step")
        if self._peer is not None:
            import ctypes
            import htgs_native as nat
            pm, dev, lib = self._peer, self.device, nat.lib()
            with torch.cuda.device(dev):
                main = torch.cuda.current_stream(dev)
                st = ctypes.c_void_p(main.cuda_stream)
                pm["hdl"].barrier(channel=0)  # every rank's small gradients and slots are complete
                # The two halves of the exchange touch different data and different resources -- the 11-float all-reduce
                # is bound by NVLink, the SH rebuild by HBM -- so they run side by side: the rebuild on a second stream,
                # forked after the barrier and joined before the closing one.
                if self._side is None:
                    self._side = torch.cuda.Stream(device=dev)
                    self._fork, self._join = torch.cuda.Event(), torch.cuda.Event()
                self._fork.record(main)
                if self.collective == "multimem":
                    bases = [self.buf.data_ptr()] * self.world  # pushed: every slot is in the local buffer
                else:
                    bases = pm["ptrs"]  # pulled: slot v is read from its owner over NVLink
                slots = [bases[v // self.views_per_rank] + 4 * self._slot_offset(v) for v in range(self.views_total)]
                self._side.wait_event(self._fork)
                with torch.cuda.stream(self._side):
                    self._expand(slots, means3D, sh_degree)
                    self._join.record(self._side)
                if self.collective == "multimem":
                    nat.check(lib.htgs_allreduce_multimem(ctypes.c_void_p(pm["mc"]), self.rank, self.world, self.n_small_pad, st),
                              "htgs_allreduce_multimem")
                else:
                    nat.check(lib.htgs_allreduce_p2p(pm["table"], self.rank, self.world, self.n_small_pad, st), "htgs_allreduce_p2p")
                main.wait_event(self._join)
                pm["hdl"].barrier(channel=0)  # nobody overwrites a slot / the flat buffer while a peer still reads it
        else:
            if self.world > 1:  # torch.distributed transport (NCCL on GPUs; gloo in the CPU tests of this host logic)
                dist.all_reduce(self.buf[:self.n_small_pad], op=dist.ReduceOp.SUM)
                lo = self._slot_offset(self.rank * self.views_per_rank)
                mine = self.buf[lo:lo + self.views_per_rank * self.slot_floats].clone()
                dist.all_gather_into_tensor(self.buf[self.n_small_pad:], mine)
            self._expand([self.buf.data_ptr() + 4 * self._slot_offset(v) for v in range(self.views_total)], means3D, sh_degree)
        return self.flat, self.sh_grad

    def _expand(self, slot_addresses, means3D, sh_degree):
        """sh_grad = sum over the view slots (device addresses, view order) of basis (x) colour gradient: htgs_sh_expand."""
        import ctypes
        import htgs_native as nat
        if not self.sh_grad.is_cuda:
            raise RuntimeError("CompactSHGradReducer: the SH rebuild is a CUDA kernel; there is no host path")
        dev = self.device
        with torch.cuda.device(dev):
            arr = (ctypes.c_void_p * self.views_total)(*slot_addresses)
            nat.check(nat.lib().htgs_sh_expand(self.P, int(sh_degree), self.M, self.views_total, means3D.detach().contiguous().data_ptr(), arr,
                                               self.sh_grad.data_ptr(), ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)),
                      "htgs_sh_expand")

    def grads(self, like):
        out = {name: self.flat[lo:hi].view_as(like[name]) for name, (lo, hi) in self.slices.items()}
        out["shs"] = self.sh_grad.view_as(like["shs"])
        return out

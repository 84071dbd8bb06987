"""Multi-GPU plumbing for the two ways the CF-3DGS optimisation shards.

One process per GPU, torch.distributed (NCCL on the B200 box, gloo in CPU tests).  The rasterizer for one view is
never split ("replicas only" inside a view); across the optimisation:
  (i)  independent frame pairs: pair i -> rank i mod world, no data-path collective while fitting, then ONE
       all-gather of the 4x4 relative poses and a sequential prefix product
       (trainer/cf3dgs_trainer.py:225-297: pose = rel_pose @ get_RT(fidx-1));
  (ii) batched global views: every rank renders its own view of a replicated model, then the optimiser step runs on
       the sum of the ranks' gradients (xyz 3 + f_dc 3 + f_rest 45 + opacity 1 + scaling 3 + rotation 4 = 59 floats
       per Gaussian, scene/gaussian_model_cf.py:254-266; 14 at SH degree 0).  Two forms of that step
       (BatchedViewsStep): shared over NVSwitch multicast -- each rank sums the replicas' gradients of its 1/N of
       the rows with multimem.ld_reduce, steps them, and writes them to every replica with multimem.st
       (gsr_gaussian_adam_step_shared on torch symmetric memory) -- or NCCL collectives (all-reduce; at SH degree >= 2
       reduce-scatter / Adam on 1/N of the SH rows / all-gather).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

GRAD_ORDER = ("means3D", "shs", "opacities", "scales", "rotations")


# ---- batched global views: layout of the flat gradient buffer and the pieces it is all-reduced in --------------------
def flat_layout(P: int, M_grad: int) -> Dict[str, tuple]:
    """Offsets (in floats) and shapes of the gradient slices inside the ONE flat buffer the rasterizer backward writes
    and NCCL reduces: rotations first (the model step reads them as float4, so their slice starts 16-byte aligned for
    any P), then means3D, opacity, scales -- the 11 floats per Gaussian of the geometry block -- then the SH gradient
    packed to its M_grad active coefficients.  (11 + 3 M_grad) P floats in all: 14 P at SH degree 0, 59 P at degree 3
    (scene/gaussian_model_cf.py:254-266 lists the parameter groups these belong to)."""
    out, off = {}, 0
    for name, shape in (("rotations", (P, 4)), ("means3D", (P, 3)), ("opacity", (P, 1)), ("scales", (P, 3)),
                        ("sh", (P, M_grad, 3))):
        if name == "sh":
            off = _sh_offset(P)      # 16-byte aligned for any P: the kernels move whole SH rows as float4 when they can
        n = 1
        for d in shape:
            n *= d
        out[name] = (off, shape)
        off += n
    out["_geometry_floats"] = (11 * P, ())
    out["_total_floats"] = (off, ())
    return out


def _sh_offset(P: int) -> int:
    return (11 * P + 3) // 4 * 4


def comm_pieces(P: int, M_grad: int, sh_chunks: int) -> List[tuple]:
    """(float offset, float count, what, row0, rows) of the pieces the flat gradient is all-reduced in, in issue order:
    the geometry block, then the SH rows in `sh_chunks` row ranges.  The model step for a piece (`what`, rows) can run
    as soon as that piece has arrived, while the next one is still on the wire."""
    pieces = [(0, 11 * P, 1, 0, 0)]
    sh_chunks = max(1, min(int(sh_chunks), P)) if P > 0 else 1
    per = -(-P // sh_chunks)
    r0 = 0
    while r0 < P:
        rows = min(per, P - r0)
        pieces.append((_sh_offset(P) + 3 * M_grad * r0, 3 * M_grad * rows, 2, r0, rows))
        r0 += rows
    return pieces


def params_checksum(tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    """Order-independent integer checksum of float32 tensors (sum of their bit patterns as int64): equal parameters give
    equal checksums on every rank, any single differing bit changes it."""
    acc = torch.zeros((), dtype=torch.int64, device=tensors[0].device)
    for t in tensors:
        acc = acc + t.detach().contiguous().view(torch.int32).to(torch.int64).sum()
    return acc


def assert_replicas_identical(tensors: Sequence[torch.Tensor]) -> bool:
    """True when every rank holds bit-identical copies (compares the checksum's max and min over the ranks)."""
    c = params_checksum(tensors)
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return True
    hi, lo = c.clone(), c.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    return bool(hi == lo)


class BatchedViewsStep:
    """One optimiser step of the batched-global-views form: every rank holds a replica of the
    global Gaussian model (gsr_model.FusedGaussianModel, device_step=True), renders ITS view, and the gradients of all
    views are summed before the same Adam step reaches every replica.

    (What follows describes the NCCL form; the multicast form is described at _setup_multicast.)
    * the rasterizer backward writes its gradients straight into slices of one flat buffer (no packing copies);
    * only the active SH bands travel (the packed gradient, dsh_M = (degree+1)^2): 14 floats per Gaussian at degree 0
      instead of 59;
    * collectives run on NCCL's own stream; the compute stream waits for each result where it needs it, the host never;
    * small SH block (degrees 0 and 1): ONE all-reduce of the whole buffer, then the model step.  Measured on 8 B200s,
      NCCL 2.28 with NVLS: all-reduces below ~100 MB run at ~400 GB/s bus bandwidth and cost 0.2 ms whatever their
      size, above it ~650 GB/s -- cutting a buffer into pieces only multiplies the fixed part;
    * large SH block (degrees 2 and 3: 4/5 of the bytes): all-reduce of the 11-float geometry block, and for the SH rows
      reduce-scatter -> Adam on this rank's 1/N of the rows -> all-gather of the updated coefficients.  Same bytes on the
      wire as an all-reduce, but the SH part of the optimiser (the larger part) runs on 1/N of the rows instead of being
      repeated on every rank, and the geometry part of the step overlaps the reduce-scatter;
    * the densification statistics are per view and stay local (sum / sum / max over ranks when they are read:
      allreduce_densify_stats).
    Every row is updated from the same reduced gradient on exactly one rank (SH) or identically on all (geometry), so the
    replicas stay bit-identical (assert_replicas_identical).  The sum over views is not divided by the number of ranks:
    Adam's update is invariant to the gradient's scale (up to eps = 1e-15).  With world size 1 the same code runs without
    a collective."""

    SHARD_FROM_M = 9       # SH coefficient count from which the SH rows are reduce-scattered / all-gathered

    def __init__(self, model, shard_features: Optional[bool] = None, multicast: Optional[bool] = False):
        """multicast: True = the shared step over NVSwitch multicast (see _setup_multicast; raises if the box cannot),
        None = use it when it can be set up on every rank and fall back to NCCL otherwise (`multicast_reason` says why),
        False = the NCCL form."""
        self.model = model
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.shard_features = shard_features
        self._key = None
        self._mc = None
        self.multicast_reason = "not requested"
        m = model
        if multicast is not False and self.world > 1:
            ok, why = True, ""
            try:
                self._setup_multicast()
            except Exception as ex:      # noqa: BLE001 -- whatever the reason, every rank has to take the same path
                ok, why = False, f"{type(ex).__name__}: {ex}"
            flag = torch.tensor([1 if ok else 0], device=m.xyz.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag) == 1:
                self.multicast_reason = "in use"
                self.rows_per, self.P_pad, self._feat_store = self._mc["rows_per"], self._mc["Pp"], None
                return
            if ok:
                self._release_multicast()
                why = "another rank could not set it up"
            self.multicast_reason = why
            if multicast is True:
                raise RuntimeError("NVSwitch multicast is not available for the shared step: " + why)
        # rows padded to a multiple of the world size so that the SH rows split evenly; the model's coefficient tensor
        # becomes a view of the padded storage (the all-gather writes straight into it)
        self.rows_per = -(-m.P // self.world)
        self.P_pad = self.rows_per * self.world
        if self.world > 1 and m.features.untyped_storage().nbytes() < self.P_pad * m.M * 3 * 4:
            store = torch.zeros((self.P_pad, m.M, 3), dtype=torch.float32, device=m.xyz.device)
            store[:m.P].copy_(m.features)
            m.features = store[:m.P]
            self._feat_store = store
        else:
            self._feat_store = None

    # ---- the shared step over NVSwitch multicast -----------------------------------------------------------------------
    # Parameters and gradients of every replica live in ONE symmetric allocation per rank (torch.distributed.
    # _symmetric_memory: same size and layout everywhere, the allocations of all ranks bound to one multicast object).
    # The rasterizer reads and writes its own copy through ordinary pointers; gsr_gaussian_adam_step_shared reads the
    # gradients of its rows through the multicast address with multimem.ld_reduce (the switch adds the replicas) and
    # writes the new parameters through it with multimem.st (the switch delivers them to every replica).  No NCCL call,
    # no reduced gradient buffer, and Adam runs on 1/N of the rows for every parameter group.
    _WIDTHS = (("rot_raw", 4), ("xyz", 3), ("opacity_raw", 1), ("log_scales", 3))

    def _setup_multicast(self):
        import torch.distributed._symmetric_memory as symm
        from diff_gaussian_rasterization._cabi import check, lib
        m, W = self.model, self.world
        dev = m.xyz.device
        rows_per = -(-(-(-m.P // W)) // 4) * 4            # rows per rank, a multiple of 4: every range moves as float4
        Pp = rows_per * W
        m._leave_dc()
        per = 11 + 3 * m.M + 3                             # floats per Gaussian: parameters (+ the dense coefficient-0 plane
                                                           # of SH degree 0, gsr_model.FusedGaussianModel.shs), and again gradients
        need = 2 * per * Pp
        # The symmetric allocation is kept and reused when the Gaussian count changes (densification rebuilds the model
        # every 100 iterations; allocating and binding a new multicast object costs ~0.1 s): it is made with 30 % headroom
        # and replaced only when a model outgrows it.  Every rank takes the same decision (same P, same history).
        key = (dev.index, W)
        cached = _SYMM_CACHE.get(key)
        if cached is not None and cached["floats"] >= need:
            whole, hdl, mc_base = cached["buf"], cached["hdl"], cached["mc0"]
        else:
            _SYMM_CACHE.pop(key, None)
            cached = None
            whole = symm.empty(int(need * 1.3) // 4 * 4, dtype=torch.float32, device=dev)
            hdl = symm.rendezvous(whole, dist.group.WORLD)
            mc_base = int(hdl.multicast_ptr)
            if mc_base == 0:
                raise RuntimeError("the symmetric allocation has no multicast address on this system")
            mc_base += whole.data_ptr() - int(hdl.buffer_ptrs[hdl.rank])
        buf, mc0 = whole[:need], mc_base
        buf.zero_()
        if cached is None:
            # self-check of the address arithmetic and of the switch: every rank writes rank + 1, the multicast read must
            # return 1 + 2 + ... + W on every rank
            probe_n = 4 * 64
            buf[:probe_n].fill_(float(self.rank + 1))
            got = torch.empty(probe_n, dtype=torch.float32, device=dev)
            hdl.barrier(channel=0)
            check(lib.gsr_multicast_probe(mc0, got.data_ptr(), probe_n, torch.cuda.current_stream().cuda_stream))
            hdl.barrier(channel=0)
            want = float(W * (W + 1) // 2)
            if not bool((got == want).all()):
                raise RuntimeError(f"multicast probe read {float(got[0])}, expected {want}")
            buf[:probe_n].zero_()
            _SYMM_CACHE[key] = dict(buf=whole, hdl=hdl, mc0=mc_base, floats=whole.numel())
        off, par, grd = 0, {}, {}
        for name, w in self._WIDTHS + (("features", 3 * m.M), ("features_dc", 3)):
            par[name] = (off, w)
            off += w * Pp
        # re-home the parameters: the model's tensors become views of the first P rows of the symmetric arrays
        old_state = m.state
        o_dc = par["features_dc"][0]
        m._dc_storage = buf[o_dc:o_dc + 3 * Pp].view(Pp, 1, 3)[:m.P]      # where shs(0) will keep coefficient 0
        for name, (o, w) in par.items():
            if name == "features_dc":
                continue
            src = getattr(m, name)
            dst = buf[o:o + w * Pp].view((Pp,) + tuple(src.shape[1:]))
            dst[:m.P].copy_(src)
            setattr(m, name, dst[:m.P])
        state = torch.zeros(lib.gsr_gaussian_adam_state_floats(Pp, m.M), dtype=torch.float32, device=dev)
        so, sn = 0, 0
        for w in (4, 4, 3, 3, 1, 1, 3, 3, 3, 3, 3 * m.M, 3 * m.M):       # the moment arrays, from P rows each to P_pad rows each
            state[sn:sn + w * m.P].copy_(old_state[so:so + w * m.P])
            so += w * m.P
            sn += w * Pp
        m.state = state
        m.shared_step_owner = "BatchedViewsStep"                   # FusedGaussianModel.step()/step_part() refuse from now on (a name, not the object: no reference cycle keeps the symmetric allocation alive)
        self._mc = dict(hdl=hdl, buf=buf, mc0=mc0, par=par, per=per, Pp=Pp, rows_per=rows_per, M_grad=None,
                        m2d=torch.empty((m.P, 3), dtype=torch.float32, device=dev))
        hdl.barrier(channel=0)     # nobody's first shared step writes into a buffer that is still being filled

    def _release_multicast(self):
        """Back to ordinary tensors (used when another rank could not set the shared step up)."""
        m, mc = self.model, self._mc
        if mc is None:
            return
        Pp, P = mc["Pp"], m.P
        m._leave_dc()
        m._dc_storage = None
        for name in mc["par"]:
            if name != "features_dc":
                setattr(m, name, getattr(m, name).clone())
        state = torch.zeros(_state_floats(P, m.M), dtype=torch.float32, device=m.xyz.device)
        so, sn = 0, 0
        for w in (4, 4, 3, 3, 1, 1, 3, 3, 3, 3, 3 * m.M, 3 * m.M):
            state[sn:sn + w * P].copy_(m.state[so:so + w * P])
            so += w * Pp
            sn += w * P
        m.state = state
        m.shared_step_owner = None
        self._mc = None

    def _multicast_step(self, settings, target, loss_fn, window):
        import diff_gaussian_rasterization as D
        m, mc = self.model, self._mc
        Pp, per, buf, P = mc["Pp"], mc["per"], mc["buf"], m.P
        deg = int(settings.sh_degree)
        M_grad = (deg + 1) * (deg + 1)
        g0 = per * Pp                                              # gradients: same order and offsets as the parameters
        if mc["M_grad"] != M_grad:                                 # the packing of the SH gradient changes with the degree
            buf[g0 + 11 * Pp:].zero_()
            mc["M_grad"] = M_grad
        feat = m.shs(deg)            # degree 0: the dense coefficient-0 plane, which lives in the symmetric buffer too
        sh_key = "features_dc" if m.dc_mode else "features"
        gv = {}
        for name, (o, w) in mc["par"].items():
            if name in ("features", "features_dc"):
                if name == sh_key:
                    gv["sh"] = buf[g0 + o:g0 + o + 3 * M_grad * P].view(P, M_grad, 3)
            else:
                gv[name] = buf[g0 + o:g0 + o + w * P].view(P, w)
        act = m.activated()
        e = torch.Tensor([])
        rs = settings
        R, color, depth, alpha, radii, geom, binning, img = D._forward_impl(
            rs.bg, m.xyz, e, act["opacities"], act["scales"], act["rotations"], rs.scale_modifier, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, feat, deg, rs.campos, False, False)
        leaf = color.requires_grad_(True)
        loss = loss_fn(leaf, target, window)
        loss.backward()
        D._backward_impl(rs.bg, m.xyz, radii, e, act["scales"], act["rotations"], rs.scale_modifier, e, rs.viewmatrix,
                         rs.projmatrix, rs.tanfovx, rs.tanfovy, leaf.grad, None, None, feat, deg, rs.campos, geom, R,
                         binning, img, alpha, False,
                         out=dict(means3D=gv["xyz"], means2D=mc["m2d"], opacity=gv["opacity_raw"], sh=gv["sh"],
                                  scales=gv["log_scales"], rotations=gv["rot_raw"]), dsh_M=M_grad)
        self._shared_update(M_grad, radii)
        return loss.detach()

    def _shared_update(self, M_grad: int, radii):
        """barrier, gsr_gaussian_adam_step_shared on this rank's rows (+ the local statistics), barrier."""
        import ctypes
        from diff_gaussian_rasterization._cabi import check, lib
        from gsr_model import LR_ORDER
        m, mc = self.model, self._mc
        Pp, per, buf, P = mc["Pp"], mc["per"], mc["buf"], m.P
        g0 = per * Pp
        m.sh_ever = max(m.sh_ever, min(M_grad, m.M))
        m.step_count += 1
        lr = (ctypes.c_float * 6)(*[float(m.lrs[k]) for k in LR_ORDER])
        local, mcast = buf.data_ptr(), mc["mc0"]
        # in dc_mode the SH tensor of the call is the dense coefficient-0 plane: a one-coefficient model (M = 1) whose
        # moments are the dense dc arrays the state buffer keeps in front of the other SH moments
        M_call = 1 if m.dc_mode else m.M
        order = ("xyz", "features_dc" if m.dc_mode else "features", "opacity_raw", "log_scales", "rot_raw")
        loc = [local + 4 * mc["par"][k][0] for k in order]
        allp = [mcast + 4 * mc["par"][k][0] for k in order]
        gsum = [mcast + 4 * (g0 + mc["par"][k][0]) for k in order]
        stream = torch.cuda.current_stream().cuda_stream
        mc["hdl"].barrier(channel=0)        # every replica's gradients are written, and nobody still reads the parameters
        with torch.cuda.device(m.xyz.device):
            check(lib.gsr_gaussian_adam_step_shared(
                Pp, M_call, min(M_grad, M_call), min(int(m.sh_ever), M_call), self.rank * mc["rows_per"], mc["rows_per"], 1, 1,
                m.step_state.data_ptr(),
                float(m.betas[0]), float(m.betas[1]), float(m.eps), lr, *loc, *allp, *gsum, m.state.data_ptr(), stream))
            check(lib.gsr_densify_stats(P, mc["m2d"].data_ptr(), radii.data_ptr(), m.xyz_gradient_accum.data_ptr(),
                                        m.denom.data_ptr(), m.max_radii2D.data_ptr(), stream))
        mc["hdl"].barrier(channel=0)        # every replica's rows have arrived here, and mine everywhere

    def buffers(self, M_grad: int):
        m = self.model
        key = (m.P, M_grad)
        if self._key != key:
            lay = flat_layout(m.P, M_grad)
            sh_off = lay["sh"][0]
            flat = torch.zeros(sh_off + self.P_pad * M_grad * 3, dtype=torch.float32, device=m.xyz.device)
            views = {}
            for name in ("rotations", "means3D", "opacity", "scales", "sh"):
                off, shape = lay[name]
                n = 1
                for d in shape:
                    n *= d
                views[name] = flat[off:off + n].view(shape)
            self._flat, self._views, self._key = flat, views, key
            self._sh_pad = flat[sh_off:]                                     # (P_pad * 3 M_grad) floats, pad rows stay zero
            self._used = lay["_total_floats"][0]
            self._m2d = torch.empty((m.P, 3), dtype=torch.float32, device=m.xyz.device)
        return self._flat, self._views

    def __call__(self, settings, target, loss_fn, window=None):
        """settings: GaussianRasterizationSettings of this rank's view; target: its image; returns this rank's loss."""
        import diff_gaussian_rasterization as D
        if self._mc is not None:
            return self._multicast_step(settings, target, loss_fn, window)
        m = self.model
        if m.P > self.P_pad or -(-m.P // self.world) != self.rows_per:
            raise RuntimeError("the model's Gaussian count changed: build a new BatchedViewsStep for it")
        deg = int(settings.sh_degree)
        M_grad = (deg + 1) * (deg + 1)
        flat, g = self.buffers(M_grad)
        act = m.activated()
        multi = self.world > 1
        shard = multi and (self.shard_features if self.shard_features is not None else M_grad >= self.SHARD_FROM_M) \
            and self._feat_store is not None
        if shard:
            m._leave_dc()          # the all-gather below moves rows of the full tensor
            feat = m.features
        else:
            feat = m.shs(deg)      # degree 0: the dense coefficient-0 plane (gsr_model.FusedGaussianModel.shs)
        e = torch.Tensor([])
        rs = settings
        R, color, depth, alpha, radii, geom, binning, img = D._forward_impl(
            rs.bg, m.xyz, e, act["opacities"], act["scales"], act["rotations"], rs.scale_modifier, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, feat, deg, rs.campos, False, False)
        leaf = color.requires_grad_(True)
        loss = loss_fn(leaf, target, window)
        loss.backward()
        D._backward_impl(rs.bg, m.xyz, radii, e, act["scales"], act["rotations"], rs.scale_modifier, e, rs.viewmatrix,
                         rs.projmatrix, rs.tanfovx, rs.tanfovy, leaf.grad, None, None, feat, deg, rs.campos, geom, R,
                         binning, img, alpha, False,
                         out=dict(means3D=g["means3D"], means2D=self._m2d, opacity=g["opacity"], sh=g["sh"],
                                  scales=g["scales"], rotations=g["rotations"]), dsh_M=M_grad)
        if not shard:
            if multi:
                dist.all_reduce(flat[:self._used])          # stream-ordered: the step below waits on the device, not the host
            m.step_part(3, g, M_grad, tick=True, d_means2D=self._m2d, radii=radii)
            return loss.detach()
        n_own = self.rows_per * 3 * M_grad
        own_grad = self._sh_pad[self.rank * n_own:(self.rank + 1) * n_own]
        w_geo = dist.all_reduce(flat[:11 * m.P], async_op=True)
        w_rs = dist.reduce_scatter_tensor(own_grad, self._sh_pad, async_op=True)       # in place: my rows of the summed SH gradient
        w_geo.wait()
        m.step_part(1, g, M_grad, tick=True, d_means2D=self._m2d, radii=radii)        # overlaps the reduce-scatter
        w_rs.wait()
        row0 = self.rank * self.rows_per
        rows = max(0, min(self.rows_per, m.P - row0))
        if rows:
            m.step_part(2, g, M_grad, row0=row0, rows=rows)
        store = self._feat_store.view(-1)
        n_feat = self.rows_per * m.M * 3
        dist.all_gather_into_tensor(store, store[self.rank * n_feat:(self.rank + 1) * n_feat])   # everyone's updated rows
        return loss.detach()


_SYMM_CACHE: Dict[tuple, dict] = {}     # (device index, world size) -> the symmetric allocation of the shared step


def release_symmetric_memory() -> None:
    """Drop the cached symmetric allocation (call on every rank, with no BatchedViewsStep in use)."""
    _SYMM_CACHE.clear()


def _state_floats(P: int, M: int) -> int:
    return 2 * (14 + 3 * M) * max(P, 0)


def shard_pairs(n_pairs: int, rank: int, world: int) -> List[int]:
    """Frame pairs (i-1, i) owned by this rank: round-robin so early and late frames mix."""
    return [i for i in range(n_pairs) if i % world == rank]


def flat_size(grads: Dict[str, torch.Tensor]) -> int:
    return sum(grads[k].numel() for k in GRAD_ORDER if k in grads and grads[k] is not None)


def flatten_into(grads: Dict[str, torch.Tensor], flat: torch.Tensor) -> torch.Tensor:
    off = 0
    for k in GRAD_ORDER:
        g = grads.get(k)
        if g is None:
            continue
        n = g.numel()
        flat[off:off + n].copy_(g.reshape(-1))
        off += n
    return flat[:off]


def unflatten_from(flat: torch.Tensor, like: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    out, off = {}, 0
    for k in GRAD_ORDER:
        g = like.get(k)
        if g is None:
            continue
        n = g.numel()
        out[k] = flat[off:off + n].view_as(g)
        off += n
    return out


def allreduce_gradients(grads: Dict[str, torch.Tensor], flat: torch.Tensor, average: bool = False):
    """Sum (or average) the per-rank gradients of a replicated Gaussian model with one collective."""
    used = flatten_into(grads, flat)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(used)
        if average:
            used.div_(dist.get_world_size())
    return unflatten_from(used, grads)


def allreduce_densify_stats(grad_accum: torch.Tensor, denom: torch.Tensor, max_radii2D: torch.Tensor):
    """Densification statistics are (sum, sum, max) across ranks (scene/gaussian_model_cf.py:716-719,
    trainer/cf3dgs_trainer.py:147-152)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(grad_accum)
        dist.all_reduce(denom)
        dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX)
    return grad_accum, denom, max_radii2D


def gather_relative_poses(local: Dict[int, torch.Tensor], n_pairs: int) -> List[torch.Tensor]:
    """All-gather the 4x4 relative poses each rank fitted; returns them indexed by pair."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    dev = next(iter(local.values())).device if local else torch.device("cpu")
    buf = torch.zeros(n_pairs, 4, 4, device=dev)
    for i, T in local.items():
        buf[i] = T
    if world > 1:
        dist.all_reduce(buf)   # disjoint ownership: a sum is a gather
    return [buf[i] for i in range(n_pairs)]


def compose_poses(rel: Sequence[torch.Tensor], first: torch.Tensor | None = None) -> List[torch.Tensor]:
    """Absolute poses by the reference's recurrence pose_i = rel_i @ pose_{i-1}
    (trainer/cf3dgs_trainer.py:294-297)."""
    cur = torch.eye(4, device=rel[0].device, dtype=rel[0].dtype) if first is None else first
    out = [cur]
    for T in rel:
        cur = T @ cur
        out.append(cur)
    return out

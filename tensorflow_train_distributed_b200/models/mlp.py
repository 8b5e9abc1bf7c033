"""The reference MLP (and its wide variant) as a fixed sequence of sm_100a kernel launches.

Model of ``forward_propagation`` / ``loss`` / ``tower_loss`` (distribute_training.py:39-110):
``[dense -> ELU -> dropout] x (L-1) -> dense`` -> mean sparse softmax cross-entropy + top-1 accuracy, with the
backward pass that ``opt.minimize`` (:152) would have built. One training step is

    prep_input                         fp32 batch -> bf16 + transposed bf16
    L x   gemm(FWD)                    the last one is FWD_XENT: loss, accuracy and dlogits in its epilogue
    L x   gemm(WGRAD [+ DGRAD])        one launch per layer, two problems per launch
    apply_update                       the fused gradient push (parameter store)

i.e. 2L + 2 launches (12 for the reference model), captured once into a CUDA graph by the trainer.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..ops import _native as N
from ..ops import gemm
from ..parallel.param_store import GpuParamStore
from ..parallel.sharding import LayerGroup, ParamLayout, mlp_groups


def _pad8(n: int) -> int:
    return -(-n // 8) * 8


def _pad4(n: int) -> int:
    return -(-n // 4) * 4


TAIL_MAX_WIDTH = 256   # layers wider than this go through the tcgen05 GEMM, not the fused tail
TAIL_MAX_BATCH = 256


def chain_forced(chain) -> bool:
    """``chain="force"`` keeps the single-CTA chain kernel although the column kernel would apply (tests, A/B runs)."""
    return chain == "force"


def chain_eligible(sizes: Sequence[int], batch: int) -> bool:
    """sizes = [n_0 .. n_L]: every layer behind the first must fit one tcgen05 tile per GEMM (K, N <= 256)."""
    return len(sizes) >= 3 and batch <= 256 and max(sizes[1:]) <= 256 and 3 * (len(sizes) - 2) <= 32


def tail_groups(layer_sizes: Sequence[int], batch: int, first_group: int = 0, enabled: bool = True):
    """Indices (into the layout's groups) of the dense layers the fused tail kernel takes over: every layer after
    the first one of the stack, provided all of them are narrow (their operands then live in L1/L2 and the chain
    is launch-latency bound). Returns () when the tail does not apply."""
    sizes = list(layer_sizes)
    n = len(sizes)
    if not enabled or n < 2 or n - 1 > N.TAIL_MAX_LAYERS or batch > TAIL_MAX_BATCH or max(sizes) > TAIL_MAX_WIDTH:
        return ()
    return tuple(range(first_group + 1, first_group + n))


class GpuMlp:
    def __init__(self, store: GpuParamStore, batch: int, *, activation: str = "elu", dropout_rate: float = 0.01,
                 seed: int = 0, pdl: bool = False, keep_logits: bool = False, first_group: int = 0,
                 feed_input: bool = True, input_grad: bool = False, chain: bool = True, net: bool = False,
                 column: bool = True):
        """``first_group`` > 0 makes this the dense head of a larger model (LeNet): its input activations
        ``a[0]`` / ``aT[0]`` are then written by the stage in front (``feed_input=False``) and, with
        ``input_grad``, the backward pass also produces ``dz0`` = dL/d(input) * act'(input) for that stage."""
        self.store, self.B = store, batch
        self.g0 = first_group
        self.feed_input, self.input_grad = feed_input, input_grad
        self.groups: List[LayerGroup] = store.layout.groups[first_group:]
        self.L = len(self.groups)
        self.act = N.ACT_CODES[activation]
        self.drop = float(dropout_rate)
        self.seed, self.pdl = seed, pdl
        dev = store.device
        B, ldB = batch, _pad8(batch)
        sizes = [self.groups[0].rows] + [g.cols for g in self.groups]  # n_0 (input) .. n_L (classes)
        self.sizes = sizes
        bf = dict(device=dev, dtype=torch.bfloat16)
        # static inputs
        self.x = torch.zeros(B, sizes[0], device=dev, dtype=torch.float32)
        self.labels = torch.zeros(B, device=dev, dtype=torch.int64)
        # activations a_0..a_{L-1} (row-major, padded) and their transposes with the ones row for the bias grad
        self.a = [torch.zeros(B, _pad8(sizes[l]), **bf) for l in range(self.L)]
        self.aT = [torch.zeros(sizes[l] + 1, ldB, **bf) for l in range(self.L)]
        for t, n in zip(self.aT, sizes):
            t[n, :B] = 1.0
        # dZ_1..dZ_L and transposes (index l-1)
        self.dz = [torch.zeros(B, _pad8(sizes[l]), **bf) for l in range(1, self.L + 1)]
        self.dzT = [torch.zeros(sizes[l], ldB, **bf) for l in range(1, self.L + 1)]
        self.dz0 = torch.zeros(B, _pad8(sizes[0]), **bf) if input_grad else None
        self.logits = torch.zeros(B, _pad8(sizes[-1]), device=dev, dtype=torch.float32) if keep_logits else None
        self.x_pool: Optional[torch.Tensor] = None
        self.labels_pool: Optional[torch.Tensor] = None
        self.use_tail = all((first_group + l) in store.tail_groups for l in range(1, self.L)) and self.L >= 2
        # chain: the narrow layers behind the first run as ONE single-CTA tcgen05 kernel (gemm_chain_kernel)
        self.use_chain = (not self.use_tail) and chain and chain_eligible(sizes, batch)
        # net: every GEMM of the stack (first layer included) in ONE launch of co-resident CTAs (gemm_net_kernel)
        self.use_net = bool(net) and not self.use_tail and not (store.pull_fused_first and first_group == 0)
        if self.use_net:
            self.use_chain = False
        # column: forward + loss + data gradients as ONE launch with the activations resident in shared memory
        # (mlp_column_kernel), then every weight gradient as one launch (net kernel) — with one replica the
        # optimiser step rides in that launch's epilogue. The default for narrow stacks.
        self.use_column = bool(column and first_group == 0 and feed_input and not input_grad and self.L >= 2
                               and not self.use_tail and not self.use_net and not chain_forced(chain)
                               and not store.pull_fused_first and gemm.column_fits(sizes[0], sizes[1:]))
        self.fused_update = self.use_column and store.world == 1
        self._column: Optional[gemm.MlpColumn] = None
        self._wnet: Optional[gemm.GemmNet] = None
        self._wnet_grads: Optional[gemm.GemmNet] = None
        self._pool_static = False
        self.x16: Optional[torch.Tensor] = None   # bf16 [P*B, ld0]      input rows of every pool slot
        self.xT16: Optional[torch.Tensor] = None  # bf16 [P*(F+1), ldB]  the same, transposed, ones row last
        if self.use_column:
            self.use_chain = False
        self._build_descs()
        if self.use_net:
            self._build_net()
        if self.use_tail:
            self._build_tail()
        if self.use_chain:
            stages = list(self.fwd_descs[1:])
            stages += [grp[1] for grp in self.bwd_descs[:-1]]   # dgrad of layers L..2: the critical path
            stages += [grp[0] for grp in self.bwd_descs[:-1]]   # then their weight gradients
            self._chain = gemm.GemmChain(stages, dev)

    def set_device_dataset(self, x_pool: Optional[torch.Tensor], labels_pool: Optional[torch.Tensor]) -> None:
        """Train from a device-resident pool [P, B, F] / [P, B]: batch ``local_step % P`` is picked on the
        device, so a captured step graph needs no host feed. ``None`` returns to the fed batch."""
        if x_pool is not None:
            assert x_pool.dim() == 3 and tuple(x_pool.shape[1:]) == tuple(self.x.shape) and x_pool.is_contiguous()
            assert labels_pool is not None and tuple(labels_pool.shape) == tuple(x_pool.shape[:2])
        self.x_pool, self.labels_pool = x_pool, labels_pool
        if self.use_column:
            self._build_column(static=x_pool is not None)

    # -------------------------------------------------------------------------------------------
    def set_feed_pool(self, feed_x: torch.Tensor, feed_y: torch.Tensor):
        """Host-fed batches: ``feed_x`` [S, B, F] fp32 / ``feed_y`` [S, B] are staging slots filled by H2D copies.
        Column mode: whoever fills slot s must also convert it (``prep_slot`` / the native feeder's prep hook) —
        the conversion then runs on the copy stream, off the step's critical path. Returns the conversion
        targets (x16, ld, xT16, ld) or None when the model converts inside the step."""
        self.x_pool, self.labels_pool = feed_x, feed_y
        if not self.use_column:
            return None
        self._build_column(static=False)
        return self.x16, self.x16.stride(0), self.xT16, self.xT16.stride(0)

    def prep_slot(self, slot: int) -> None:
        """fp32 slot -> bf16 rows + transposed rows of the same slot (column mode), on the current stream."""
        B, F = self.B, self.sizes[0]
        src = self.x_pool[slot] if self.x_pool is not None else self.x
        gemm.prep_input(src, self.x16[slot * B:(slot + 1) * B], self.xT16[slot * (F + 1):slot * (F + 1) + F])

    def _build_column(self, static: bool) -> None:
        s, B, sz = self.store, self.B, self.sizes
        dev = s.device
        F, ldB = sz[0], _pad8(B)
        P = int(self.x_pool.shape[0]) if self.x_pool is not None else 1
        self._pool_static = static
        if self._column is not None:
            self._column.close()
        self.x16 = torch.zeros(P * B, _pad8(F), device=dev, dtype=torch.bfloat16)
        self.xT16 = torch.zeros(P * (F + 1), ldB, device=dev, dtype=torch.bfloat16)
        self.xT16.view(P, F + 1, ldB)[:, F, :B] = 1.0  # ones row: the bias gradient falls out of the wgrad GEMM
        if static:  # a device-resident data set is converted once, here
            for slot in range(P):
                self.prep_slot(slot)
        d = N.ColumnDesc()
        d.B, d.n_layers, d.act, d.drop_rate = B, self.L, self.act, self.drop
        d.cluster, d.pool_batches = 0, P
        d.x_pool, d.ld_x = self.x16.data_ptr(), self.x16.stride(0)
        labels = self.labels_pool if self.labels_pool is not None else self.labels
        d.labels_pool = labels.data_ptr()
        d.metrics = s.metrics_acc_ptr
        d.step_ptr = s.local_step_ptr
        d.global_step_ptr = s.global_step_ptr
        d.lr_out = s.lr_slot.data_ptr()
        d.hyper = s.hyper_native
        for l in range(1, self.L + 1):
            y = d.layer[l - 1]
            gi = self.g0 + l - 1
            y.in_, y.out, y.layer_id = sz[l - 1], sz[l], l
            y.w_t, y.ld_t = s.w_t[gi].data_ptr(), s.w_t[gi].stride(0)
            if l >= 2:
                y.w_rm, y.ld_rm = s.w_rm[gi].data_ptr(), s.w_rm[gi].stride(0)
            y.bias = s.bias(gi).data_ptr()
            if l < self.L:
                y.aT, y.ld_aT = self.aT[l].data_ptr(), self.aT[l].stride(0)
            y.dzT, y.ld_dzT = self.dzT[l - 1].data_ptr(), self.dzT[l - 1].stride(0)
        self._column = gemm.MlpColumn(d)
        # every weight gradient in ONE launch; with a single replica the update happens in its epilogue
        self._wnet = self._build_wnet(fused=self.fused_update)
        self._wnet_grads = None if self.fused_update else self._wnet  # plain gradients: built on first use

    def _build_wnet(self, fused: bool) -> "gemm.GemmNet":
        s, B, sz = self.store, self.B, self.sizes
        F = sz[0]
        P = self.xT16.shape[0] // (F + 1)
        descs = []
        for l in range(self.L, 0, -1):
            gi = self.g0 + l - 1
            a = self.aT[l - 1] if l >= 2 else self.xT16
            kw = dict(a_slot_rows=F + 1, a_slots=P) if l == 1 else {}
            if fused:
                descs.append(gemm.make_desc(a, self.dzT[l - 1], sz[l - 1] + 1, sz[l], B, N.EPI_WGRAD_UPDATE, scale=1.0,
                                            out=s.w_rm[gi], outT=s.w_t[gi], upd_w=s.master_matrix(gi),
                                            upd_ema=s.ema_matrix(gi) if s.hyper.ema_decay > 0 else None,
                                            upd_bias=s.bias(gi), upd_regularize=s.layout.groups[gi].regularize, **kw))
            else:
                descs.append(gemm.make_desc(a, self.dzT[l - 1], sz[l - 1] + 1, sz[l], B, N.EPI_WGRAD,
                                            outf=s.grad_matrix(gi), scale=1.0, **kw))
        return gemm.GemmNet(descs, [[] for _ in descs], s.device)

    # -------------------------------------------------------------------------------------------
    def _build_descs(self) -> None:
        s, B, sz, g0 = self.store, self.B, self.sizes, self.g0
        self.fwd_descs, self.bwd_descs = [], []
        for l in range(1, self.L + 1):
            if self.use_tail and l >= 2:
                break  # layers 2..L belong to the fused tail kernel
            last = l == self.L
            kw = dict(bias=s.bias(g0 + l - 1), layer_id=l)
            if last:
                d = gemm.make_desc(self.a[l - 1], s.w_t[g0 + l - 1], B, sz[l], sz[l - 1], N.EPI_FWD_XENT,
                                   out=self.dz[l - 1], outT=self.dzT[l - 1], outf=self.logits, labels=self.labels,
                                   metrics_ptr=s.metrics_acc_ptr, scale=1.0, **kw)
            else:
                if l == 1 and g0 == 0:
                    ptrs, kb = s.first_layer_pull_plan()  # K3: all-gather of W1 fused into this GEMM
                    kw.update(b_peers=ptrs, k_bounds=kb)
                d = gemm.make_desc(self.a[l - 1], s.w_t[g0 + l - 1], B, sz[l], sz[l - 1], N.EPI_FWD, act=self.act,
                                   drop_rate=self.drop, out=self.a[l], outT=self.aT[l], **kw)
            self.fwd_descs.append(d)
        for l in range(self.L, 0, -1):
            if self.use_tail and l >= 2:
                continue
            grp = [gemm.make_desc(self.aT[l - 1], self.dzT[l - 1], sz[l - 1] + 1, sz[l], B, N.EPI_WGRAD,
                                  outf=s.grad_matrix(g0 + l - 1), scale=1.0)]
            if l > 1:
                grp.append(gemm.make_desc(self.dz[l - 1], s.w_rm[g0 + l - 1], B, sz[l - 1], sz[l], N.EPI_DGRAD, act=self.act,
                                          drop_rate=self.drop, layer_id=l - 1, aux=self.a[l - 1],
                                          out=self.dz[l - 2], outT=self.dzT[l - 2]))
            elif self.input_grad:  # gradient for the stage in front: times act'(its output), no dropout there
                grp.append(gemm.make_desc(self.dz[0], s.w_rm[g0], B, sz[0], sz[1], N.EPI_DGRAD, act=self.act,
                                          drop_rate=0.0, layer_id=0, aux=self.a[0], out=self.dz0))
            self.bwd_descs.append(grp)

    def _build_net(self) -> None:
        """Problem list + dependencies of the net kernel: F1..FL, then the data gradients DL..D2 (D1 when the stage
        in front wants one) — the critical path — then the weight gradients WL..W1."""
        L = self.L
        descs = list(self.fwd_descs)
        deps: List[List[int]] = [[] if l == 0 else [l - 1] for l in range(L)]
        dz_producer = {L: L - 1}  # dz_L comes out of the last forward layer's softmax-xent epilogue
        for i, grp in enumerate(self.bwd_descs):  # layers L..1
            if len(grp) > 1:
                l = L - i
                descs.append(grp[1])
                deps.append([dz_producer[l]])
                dz_producer[l - 1] = len(descs) - 1
        for i, grp in enumerate(self.bwd_descs):
            l = L - i
            descs.append(grp[0])
            deps.append([dz_producer[l]])
        self._net = gemm.GemmNet(descs, deps, self.store.device)

    # -------------------------------------------------------------------------------------------
    def _build_tail(self) -> None:
        """Descriptor + fp32 scratch of the fused tail (layers 2..L of this stack): see csrc/mlp_tail.cu."""
        s, B, sz, g0 = self.store, self.B, self.sizes, self.g0
        dev = s.device
        d = N.TailDesc()
        d.B, d.n_layers, d.cluster = B, self.L - 1, 0
        d.in_act_kind, d.in_drop_rate, d.in_layer_id = self.act, self.drop, 1
        d.a_in, d.ld_a_in = self.a[1].data_ptr(), self.a[1].stride(0)
        d.a_inT, d.ld_a_inT = self.aT[1].data_ptr(), self.aT[1].stride(0)
        d.dz_in, d.ld_dz_in = self.dz[0].data_ptr(), self.dz[0].stride(0)
        d.dz_inT, d.ld_dz_inT = self.dzT[0].data_ptr(), self.dzT[0].stride(0)
        d.labels = self.labels.data_ptr()
        d.metrics = s.metrics_acc_ptr
        d.step_ptr = s.local_step_ptr
        self._tail_bufs = []
        f32 = dict(device=dev, dtype=torch.float32)
        for j in range(self.L - 1):
            l = j + 2  # 1-based layer number inside this stack
            gi = g0 + l - 1
            y = d.layer[j]
            y.in_, y.out = sz[l - 1], sz[l]
            y.act_kind, y.drop_rate, y.layer_id = self.act, self.drop, l
            y.w, y.ld_w = s.w_f32[gi].data_ptr(), s.w_f32[gi].stride(0)
            y.wt, y.ld_wt = s.wt_f32[gi].data_ptr(), s.wt_f32[gi].stride(0)
            y.bias = s.bias(gi).data_ptr()
            y.dw = s.grad_matrix(gi).data_ptr()
            act = torch.zeros(B, _pad4(sz[l]), **f32)
            actT = torch.zeros(_pad4(sz[l] + 1), _pad4(B), **f32)
            actT[sz[l], :B] = 1.0
            dz = torch.zeros(_pad4(B), _pad4(sz[l]), **f32)
            y.act, y.ld_act = act.data_ptr(), act.stride(0)
            y.actT, y.ld_actT = actT.data_ptr(), actT.stride(0)
            y.dz, y.ld_dz = dz.data_ptr(), dz.stride(0)
            self._tail_bufs.append((act, actT, dz))
        self._tail_desc = d
        if self.logits is not None:  # the last layer's "activation" buffer holds the logits
            self.logits = self._tail_bufs[-1][0]

    def launch_tail(self) -> int:
        import ctypes as C
        lib = self.store.lib
        N.check(lib.ttd_mlp_tail(C.byref(self._tail_desc), C.c_uint64(self.seed & (2**64 - 1)), 1 if self.pdl else 0,
                                 N.cur_stream()), "ttd_mlp_tail")
        return 1

    # -------------------------------------------------------------------------------------------
    def launch_forward_backward(self) -> int:
        """Enqueue forward + backward on the current stream; returns the number of kernels launched."""
        step_ptr = self.store.local_step_ptr
        n = 0
        if self.use_column:
            if self._column is None:
                self._build_column(static=False)
            if self.x_pool is None:  # single fed batch: converted inside the step
                self.prep_slot(0)
                n += 1
            return n + self.launch_forward() + self.launch_backward()
        if not self.feed_input:
            pass
        elif self.x_pool is not None:
            gemm.prep_input(self.x_pool, self.a[0], self.aT[0], labels_pool=self.labels_pool, labels_out=self.labels,
                            step_ptr=step_ptr, pdl=self.pdl)
        else:
            gemm.prep_input(self.x, self.a[0], self.aT[0], pdl=self.pdl)
        n += 1 if self.feed_input else 0
        return n + self.launch_forward() + self.launch_backward()

    def launch_forward(self) -> int:
        step_ptr, n = self.store.local_step_ptr, 0
        if self.use_column:  # forward, loss and every data gradient: one launch
            self._column.launch(seed=self.seed, pdl=self.pdl)
            return 1
        if self.use_net:  # forward, loss and the whole backward: one launch
            self._net.launch(seed=self.seed, step_ptr=step_ptr, pdl=self.pdl)
            return 1
        if self.use_tail:  # first layer on the tensor cores, everything behind it in one fused launch
            gemm.gemm_group([self.fwd_descs[0]], seed=self.seed, step_ptr=step_ptr, pdl=self.pdl)
            return 1 + self.launch_tail()
        if self.use_chain:  # layers 2..L forward + loss + their whole backward: one single-CTA launch
            gemm.gemm_group([self.fwd_descs[0]], seed=self.seed, step_ptr=step_ptr, pdl=self.pdl)
            self._chain.launch(seed=self.seed, step_ptr=step_ptr, pdl=self.pdl)
            return 2
        for d in self.fwd_descs:
            gemm.gemm_group([d], seed=self.seed, step_ptr=step_ptr, pdl=self.pdl)
            n += 1
        return n

    def launch_backward(self, update: bool = False, bump_global_step: bool = True) -> int:
        """Weight gradients (and what is left of the data gradients). ``update`` (column mode, one replica): the
        optimiser step and the end-of-step bookkeeping ride in the same launch."""
        step_ptr, n = self.store.local_step_ptr, 0
        if self.use_column:
            s = self.store
            if update and self.fused_update:
                self._wnet.launch(seed=self.seed, step_ptr=step_ptr, pdl=self.pdl, hyper=s.hyper_native,
                                  lr_slot=s.lr_slot.data_ptr(), state_ptr=s.state_ptr,
                                  finalize=1 | (2 if bump_global_step else 0))
            else:  # plain gradients into the flat buffer (several replicas, async mode, gradient checks)
                if self._wnet_grads is None:
                    self._wnet_grads = self._build_wnet(fused=False)
                self._wnet_grads.launch(seed=self.seed, step_ptr=step_ptr, pdl=self.pdl)
            return 1
        if self.use_net:
            return 0
        if self.use_tail or self.use_chain:  # dz of the first layer and every other gradient exist already
            gemm.gemm_group(self.bwd_descs[-1], seed=self.seed, step_ptr=step_ptr, pdl=self.pdl)
            return 1
        for grp in self.bwd_descs:
            gemm.gemm_group(grp, seed=self.seed, step_ptr=step_ptr, pdl=self.pdl)
            n += 1
        return n

    def launch_step(self, bump_global_step: bool = True, sync: bool = True) -> int:
        """sync: fwd/bwd + the fused push (aggregate of all replicas). async: pull + fwd/bwd + atomic push."""
        if sync and self.use_column and self.fused_update:
            if self._column is None:
                self._build_column(static=False)
            n = 0
            if self.x_pool is None:
                self.prep_slot(0)
                n = 1
            return n + self.launch_forward() + self.launch_backward(update=True, bump_global_step=bump_global_step)
        if sync:
            n = self.launch_forward_backward()
            self.store.apply_update(bump_global_step, pdl=self.pdl)
            return n + 1
        self.store.async_pull()
        n = self.launch_forward_backward()
        self.store.async_push()
        return n + 2


def build_layout(input_dim: int, layer_sizes: Sequence[int]) -> ParamLayout:
    return ParamLayout(mlp_groups(input_dim, layer_sizes))

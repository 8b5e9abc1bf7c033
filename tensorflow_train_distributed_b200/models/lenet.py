"""LeNet-5-style CNN on 28x28x1 (BASELINE config 5; the reference has no CNN — SURVEY §0.3 [NS]).

    conv1 5x5 SAME 1->6   + act + pool 2x2   28x28 -> 14x14      (CUDA-core direct conv, csrc/conv.cu)
    conv2 5x5 VALID 6->16 + act + pool 2x2   14x14 -> 10x10 -> 5x5
    fc1 400->120, fc2 120->84 (act + dropout), output 84->10     (tcgen05 GEMM stack, shared with the MLP)

Variables under TF names with HWIO kernels: conv1/kernel [5,5,1,6], conv2/kernel [5,5,6,16], fc1/kernel
[400,120], fc2/kernel [120,84], output/kernel [84,10] (+ biases) — 61,706 parameters. The parameter store
treats a conv kernel as the matrix [KH*KW*Cin, Cout], which is exactly the HWIO tensor flattened.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from ..ops import _native as N
from ..parallel.param_store import GpuParamStore
from ..parallel.sharding import ParamLayout, lenet_groups
from .mlp import GpuMlp


def build_layout(num_classes: int = 10) -> ParamLayout:
    return ParamLayout(lenet_groups(num_classes))


class _ConvStage:
    def __init__(self, store: GpuParamStore, group: int, B: int, H: int, W: int, cin: int, cout: int, same: bool,
                 act: int):
        self.store, self.group, self.act = store, group, act
        self.desc = N.ConvDesc(B, H, W, cin, cout, 5, 5, 1 if same else 0)
        oh = H if same else H - 4
        self.PH = self.PW = oh // 2
        dev = store.device
        self.pooled = torch.zeros(B, self.PH, self.PW, cout, device=dev, dtype=torch.float32)
        self.argmax = torch.zeros(B, self.PH, self.PW, cout, device=dev, dtype=torch.uint8)
        self.w = store.w_rm[group]
        # weight-gradient slices: arrival counters + per-slice partial rows (zeroed once, re-armed by the kernel)
        k1 = 25 * cin + 1
        self.wgrad_scratch = torch.zeros(-(-k1 // 64) * 64 + 64 * k1 * 16, device=dev, dtype=torch.float32)
        assert self.w is not None, "conv stages read the TF-layout compute copy (need_w_rm_first=True)"

    def forward(self, x: torch.Tensor, flat: Optional[torch.Tensor] = None, flatT: Optional[torch.Tensor] = None) -> None:
        lib = self.store.lib
        N.check(lib.ttd_conv_pool_fwd(C.byref(self.desc), x.data_ptr(), self.w.data_ptr(), self.w.stride(0),
                                      self.store.bias(self.group).data_ptr(), self.act, self.pooled.data_ptr(),
                                      self.argmax.data_ptr(), N.ptr(flat), flat.stride(0) if flat is not None else 0,
                                      N.ptr(flatT), flatT.stride(0) if flatT is not None else 0, N.cur_stream()),
                "ttd_conv_pool_fwd")

    def backward(self, dzp: torch.Tensor, x: torch.Tensor, dx: Optional[torch.Tensor], act_below: int) -> int:
        lib = self.store.lib
        ld = dzp.stride(0) if dzp.dim() == 2 else dzp[0].numel()
        N.check(lib.ttd_conv_pool_bwd(C.byref(self.desc), dzp.data_ptr(), ld, 1 if dzp.dtype == torch.bfloat16 else 0,
                                      self.argmax.data_ptr(), x.data_ptr(), self.w.data_ptr(), self.w.stride(0),
                                      self.store.grad_matrix(self.group).data_ptr(), N.ptr(dx), x.data_ptr(),
                                      act_below, self.wgrad_scratch.data_ptr(), self.wgrad_scratch.numel(),
                                      N.cur_stream()), "ttd_conv_pool_bwd")
        return 2 if dx is not None else 1


class GpuLeNet:
    def __init__(self, store: GpuParamStore, batch: int, *, activation: str = "elu", dropout_rate: float = 0.01,
                 seed: int = 0, pdl: bool = False, chain: bool = True, net: bool = False):
        self.store, self.B = store, batch
        act = N.ACT_CODES[activation]
        dev = store.device
        self.x = torch.zeros(batch, 784, device=dev, dtype=torch.float32)  # NHWC with C = 1
        self.conv1 = _ConvStage(store, 0, batch, 28, 28, 1, 6, True, act)
        self.conv2 = _ConvStage(store, 1, batch, 14, 14, 6, 16, False, act)
        self.head = GpuMlp(store, batch, activation=activation, dropout_rate=dropout_rate, seed=seed, pdl=pdl,
                           first_group=2, feed_input=False, input_grad=True, chain=chain, net=net)
        self.labels = self.head.labels
        self.dpool1 = torch.zeros(batch, 14, 14, 6, device=dev, dtype=torch.float32)
        self.act = act
        self.x_pool = self.labels_pool = None

    def set_device_dataset(self, x_pool, labels_pool) -> None:
        """Device-resident data set [P, B, 784] / [P, B]: the batch of ``local_step % P`` is staged on the device."""
        self.x_pool, self.labels_pool = x_pool, labels_pool

    def launch_forward_backward(self) -> int:
        from ..ops import gemm
        extra = 0
        if self.x_pool is not None:
            gemm.prep_input(self.x_pool, None, None, labels_pool=self.labels_pool, labels_out=self.labels,
                            step_ptr=self.store.local_step_ptr, xf=self.x)
            extra = 1
        self.conv1.forward(self.x)
        self.conv2.forward(self.conv1.pooled, flat=self.head.a[0], flatT=self.head.aT[0])
        n = 2 + extra + self.head.launch_forward() + self.head.launch_backward()
        n += self.conv2.backward(self.head.dz0, self.conv1.pooled, self.dpool1, self.act)
        n += self.conv1.backward(self.dpool1, self.x, None, N.ACT_NONE)
        return n

    def launch_step(self, bump_global_step: bool = True, sync: bool = True) -> int:
        if sync:
            n = self.launch_forward_backward()
            self.store.apply_update(bump_global_step)
            return n + 1
        self.store.async_pull()
        n = self.launch_forward_backward()
        self.store.async_push()
        return n + 2

"""Optimizer configuration for the sparse row update (SPEC A5)."""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

from ._lib import EDL_OPT_ADAGRAD, EDL_OPT_ADAM, OptParams


@dataclass
class OptimizerSpec:
    kind: int = EDL_OPT_ADAGRAD
    lr: float = 0.01
    eps: float = 1e-7
    init_accum: float = 0.1     # Adagrad accumulator start value
    beta1: float = 0.9
    beta2: float = 0.999

    @staticmethod
    def adagrad(lr: float = 0.01, eps: float = 1e-7, init_accum: float = 0.1) -> "OptimizerSpec":
        return OptimizerSpec(EDL_OPT_ADAGRAD, lr, eps, init_accum)

    @staticmethod
    def adam(lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999,
             eps: float = 1e-8) -> "OptimizerSpec":
        return OptimizerSpec(EDL_OPT_ADAM, lr, eps, 0.0, beta1, beta2)

    @property
    def n_slots(self) -> int:
        return 1 if self.kind == EDL_OPT_ADAGRAD else 2

    @property
    def name(self) -> str:
        return "adagrad" if self.kind == EDL_OPT_ADAGRAD else "adam"

    def slot_init(self, k: int) -> float:
        return self.init_accum if self.kind == EDL_OPT_ADAGRAD else 0.0

    def rate(self, step: int) -> float:
        """The fp32 rate of global step `step` (1-based).  Adam's bias-corrected rate
        lr_t = lr*sqrt(1-b2^t)/(1-b1^t) is evaluated in float64 and rounded once."""
        if self.kind == EDL_OPT_ADAGRAD:
            lr = self.lr
        else:
            lr = self.lr * math.sqrt(1.0 - self.beta2 ** step) / (1.0 - self.beta1 ** step)
        return float(np.float32(lr))

    def params(self, step: int, lr_dev: int = 0) -> OptParams:
        """Kernel parameters for global step `step`.  lr_dev: address of a device float the kernels
        read the rate from instead (the caller stores rate(step) there; CUDA-graph replay)."""
        return OptParams(self.kind, self.rate(step), float(np.float32(self.eps)),
                         float(np.float32(self.beta1)), float(np.float32(self.beta2)), lr_dev or None)

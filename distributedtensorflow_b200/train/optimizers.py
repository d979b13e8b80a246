"""TF1 optimizers whose *apply* step runs on the parameter server that owns the variable.

Parity targets: ``tf.train.AdamOptimizer(0.0005).minimize(loss, global_step)``
(run_mnist_distributed.py:116), ``AdagradOptimizer(0.01)``
(templates/00_between_graph_replication_async_mnist.py:34), ``AdamOptimizer(learning_rate)``
(templates/00_mnist_replica.py:166) plus SGD and Momentum. Update rules are TF1's exactly
(python/training/{adam,adagrad,momentum,gradient_descent}.py):

* Adam: ``lr_t = lr*sqrt(1-b2^t)/(1-b1^t)``, ``m += (g-m)(1-b1)``, ``v += (g*g-v)(1-b2)``,
  ``var -= lr_t*m/(sqrt(v)+eps)``; ``beta1_power``/``beta2_power`` are *shared PS variables*
  multiplied by beta after every apply, so ``t`` counts applies from all workers.
* Adagrad: ``accum(init 0.1) += g*g; var -= lr*g/sqrt(accum)``.
* Momentum: ``accum = mu*accum + g; var -= lr*accum`` (Nesterov optional).
* SGD: ``var -= lr*g``.

``use_locking=False`` (the default, and what the reference uses) means Hogwild: concurrent
workers' applies interleave without mutual exclusion.

``minimize`` returns a ``TrainOp`` node; the session's executor computes gradients on the worker
and hands them to the PS transport's ``push_apply`` together with this optimizer's spec. The
CPU transports call :meth:`Optimizer.apply_dense` (PyTorch, in place); the B200 transport maps the
same spec onto the fused ``dtf_push_apply`` kernel.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence

from ..framework import graph as G

OPT_CODES = {"sgd": 0, "momentum": 1, "adagrad": 2, "adam": 3}


@dataclass
class OptimizerSpec:
    """Everything a PS shard needs to apply a gradient (picklable; sent over the control plane)."""
    kind: str
    learning_rate: float
    beta1: float = 0.9
    beta2: float = 0.999
    epsilon: float = 1e-8
    momentum: float = 0.0
    use_nesterov: bool = False
    initial_accumulator_value: float = 0.1
    use_locking: bool = False
    name: str = ""

    @property
    def code(self) -> int:
        return OPT_CODES[self.kind]

    @property
    def num_slots(self) -> int:
        return {"sgd": 0, "momentum": 1, "adagrad": 1, "adam": 2}[self.kind]

    def slot_names(self) -> List[str]:
        """TF checkpoint slot suffixes: ``<var>/Adam``, ``<var>/Adam_1`` etc."""
        n = self.name
        return {"sgd": [], "momentum": [n], "adagrad": [n], "adam": [n, n + "_1"]}[self.kind]

    def slot_init(self, slot_index: int) -> float:
        return self.initial_accumulator_value if self.kind == "adagrad" else 0.0

    def nonslot_names(self) -> List[str]:
        return ["beta1_power", "beta2_power"] if self.kind == "adam" else []


class TrainOp(G.Tensor):
    """The op ``session.run`` executes for one training step."""

    def __init__(self, optimizer: "Optimizer", loss: G.Tensor, var_list: Sequence[G.Variable],
                 global_step: Optional[G.Variable], name: str):
        super().__init__("TrainOp", [loss], {}, (), "none", name)
        self.optimizer = optimizer
        self.loss = loss
        self.var_list = list(var_list)
        self.global_step = global_step
        self.sync = None  # set by SyncReplicasOptimizer


class Optimizer:
    kind = "sgd"

    def __init__(self, learning_rate: float, use_locking: bool = False, name: str = "Optimizer"):
        self._lr = float(learning_rate)
        self._use_locking = use_locking
        self._name = name

    # -- TF API ---------------------------------------------------------------------------------
    def get_name(self) -> str:
        return self._name

    def spec(self) -> OptimizerSpec:
        return OptimizerSpec(self.kind, self._lr, use_locking=self._use_locking, name=self._name)

    def compute_gradients(self, loss: G.Tensor, var_list: Optional[Sequence[G.Variable]] = None):
        vars_ = list(var_list) if var_list is not None else loss.graph.trainable_variables()
        if not vars_:
            raise ValueError("No variables to optimize.")
        return [(G.Tensor("Gradient", [loss, v], {}, v.shape, v.dtype, f"gradients/{v.name}_grad"), v)
                for v in vars_]

    def apply_gradients(self, grads_and_vars, global_step: Optional[G.Variable] = None,
                        name: Optional[str] = None) -> TrainOp:
        grads_and_vars = list(grads_and_vars)
        if not grads_and_vars:
            raise ValueError("No variables provided.")
        loss = grads_and_vars[0][0].inputs[0]
        return TrainOp(self, loss, [v for _, v in grads_and_vars], global_step, name or self._name)

    def minimize(self, loss: G.Tensor, global_step: Optional[G.Variable] = None,
                 var_list: Optional[Sequence[G.Variable]] = None, name: Optional[str] = None) -> TrainOp:
        return self.apply_gradients(self.compute_gradients(loss, var_list), global_step, name)

    def get_slot_names(self) -> List[str]:
        return self.spec().slot_names()

    # -- PS-side apply (CPU transports) ------------------------------------------------------------
    @staticmethod
    def apply_dense(spec: OptimizerSpec, var, grad, slots: Sequence, powers: Optional[dict]) -> None:
        """In-place TF1 update of ``var`` (torch tensors). ``powers`` = {"beta1_power": tensor,
        "beta2_power": tensor} for Adam (read here, advanced by :meth:`finish`)."""
        import torch
        with torch.no_grad():
            if spec.kind == "sgd":
                var.add_(grad, alpha=-spec.learning_rate)
            elif spec.kind == "momentum":
                accum = slots[0]
                accum.mul_(spec.momentum).add_(grad)
                if spec.use_nesterov:
                    var.add_(grad, alpha=-spec.learning_rate)
                    var.add_(accum, alpha=-spec.learning_rate * spec.momentum)
                else:
                    var.add_(accum, alpha=-spec.learning_rate)
            elif spec.kind == "adagrad":
                accum = slots[0]
                accum.addcmul_(grad, grad)
                var.addcdiv_(grad, accum.sqrt(), value=-spec.learning_rate)
            elif spec.kind == "adam":
                m, v = slots
                b1p = float(powers["beta1_power"])
                b2p = float(powers["beta2_power"])
                lr_t = spec.learning_rate * math.sqrt(1.0 - b2p) / (1.0 - b1p)
                m.add_(grad - m, alpha=1.0 - spec.beta1)
                v.add_(grad * grad - v, alpha=1.0 - spec.beta2)
                var.addcdiv_(m, v.sqrt().add_(spec.epsilon), value=-lr_t)
            else:
                raise ValueError(spec.kind)

    @staticmethod
    def finish(spec: OptimizerSpec, powers: Optional[dict]) -> None:
        """TF's ``_finish``: advance the shared beta powers once per apply_gradients."""
        if spec.kind == "adam":
            powers["beta1_power"].mul_(spec.beta1)
            powers["beta2_power"].mul_(spec.beta2)


class GradientDescentOptimizer(Optimizer):
    kind = "sgd"

    def __init__(self, learning_rate, use_locking=False, name="GradientDescent"):
        super().__init__(learning_rate, use_locking, name)


class MomentumOptimizer(Optimizer):
    kind = "momentum"

    def __init__(self, learning_rate, momentum, use_locking=False, name="Momentum", use_nesterov=False):
        super().__init__(learning_rate, use_locking, name)
        self._momentum = float(momentum)
        self._use_nesterov = use_nesterov

    def spec(self) -> OptimizerSpec:
        return OptimizerSpec("momentum", self._lr, momentum=self._momentum,
                             use_nesterov=self._use_nesterov, use_locking=self._use_locking,
                             name=self._name)


class AdagradOptimizer(Optimizer):
    kind = "adagrad"

    def __init__(self, learning_rate, initial_accumulator_value=0.1, use_locking=False, name="Adagrad"):
        if initial_accumulator_value <= 0.0:
            raise ValueError("initial_accumulator_value must be positive: %s" % initial_accumulator_value)
        super().__init__(learning_rate, use_locking, name)
        self._init_acc = float(initial_accumulator_value)

    def spec(self) -> OptimizerSpec:
        return OptimizerSpec("adagrad", self._lr, initial_accumulator_value=self._init_acc,
                             use_locking=self._use_locking, name=self._name)


class AdamOptimizer(Optimizer):
    kind = "adam"

    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-8, use_locking=False,
                 name="Adam"):
        super().__init__(learning_rate, use_locking, name)
        self._beta1, self._beta2, self._epsilon = float(beta1), float(beta2), float(epsilon)

    def spec(self) -> OptimizerSpec:
        return OptimizerSpec("adam", self._lr, self._beta1, self._beta2, self._epsilon,
                             use_locking=self._use_locking, name=self._name)

"""Host-side PS replica: what runs in place of `server.join()`
(examples/workdir/mnist_replica.py:121-122) in a pod started with
`--job_name=ps --task_index=i`.

It owns task i's variables and optimizer slots in HBM as one flat arena (the
bucket that tf.train.replica_device_setter's round-robin gives task i, or the
i-th contiguous shard in flat mode), and applies each step's N worker gradients
with one fused launch through the C ABI. Semantics follow TF-1.4: SUM or MEAN
aggregation in worker order, Adam's beta powers advanced once per step
(adam.py _finish), global_step incremented once per apply.
"""
from __future__ import annotations

import torch

from . import accel, placement


class ParameterServer:
    def __init__(self, argv=None, *, task_index=None, n_ps=None, var_elems, optimizer,
                 hparams: accel.HParams, sharding="round_robin", first_task=0, align_elems=64,
                 ctx: accel.Context | None = None, gpu_ordinals=None, device_state=False):
        if argv is not None:
            job, task_index, _, n_ps = placement.parse_argv(list(argv))
            if job != "ps":
                raise ValueError(f"ParameterServer needs --job_name=ps, got {job}")
        if task_index is None or n_ps is None:
            raise ValueError("either argv or (task_index, n_ps) is required")
        self.task, self.n_ps = task_index, n_ps
        self.ctx = ctx or accel.Context(n_ps, gpu_ordinals)
        self.device = torch.device("cuda", self.ctx.gpu_of(self.task))
        self.optimizer, self.hp = optimizer, hparams
        self.global_step = 0
        var_elems = list(var_elems)
        if sharding == "round_robin":
            task, off, fill = accel.plan_round_robin(var_elems, n_ps, first_task, align_elems)
            # (variable index, offset in this task's arena, elements) for the variables held here
            self.layout = [(k, off[k], var_elems[k]) for k in range(len(var_elems))
                           if task[k] == self.task]
            self.count = fill[self.task]
            self.arena_offset = 0
        elif sharding == "flat":
            shards, _ = accel.plan_flat(sum(var_elems), n_ps, align_elems)
            self.arena_offset, self.count = shards[self.task]
            self.layout = None
        else:
            raise ValueError(f"unknown sharding {sharding!r}")
        n = max(self.count, 1)
        self.var = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.slot0 = torch.zeros(n, dtype=torch.float32, device=self.device) \
            if optimizer >= accel.OPT_MOMENTUM else None
        self.slot1 = torch.zeros(n, dtype=torch.float32, device=self.device) \
            if optimizer == accel.OPT_ADAM else None

        # optional: Adam's beta powers and the global step live on the GPU and advance there, so a
        # step has no host-side mutable input (it can be recorded once and replayed)
        self.dev_state = None
        if device_state:
            self.dev_state = torch.zeros(4, dtype=torch.float32, device=self.device)
            self.ctx.opt_state_init(self.dev_state, self.hp, 0)

    def device_global_step(self) -> int:
        return int(self.dev_state[2:4].view(torch.int64).item())

    def variable(self, k: int) -> torch.Tensor:
        """View of variable k (creation index) inside this task's arena."""
        for kk, off, n in self.layout or ():
            if kk == k:
                return self.var[off:off + n]
        raise KeyError(f"variable {k} is not placed on PS task {self.task}")

    def apply_gradients(self, worker_grads, stream=None) -> None:
        """One step: worker_grads[w] is worker w's gradient for this task's arena
        (device tensor of `count` elements, fp32 or bf16)."""
        if self.dev_state is not None:
            self.ctx.apply_dev_state(self.optimizer, self.var, self.slot0, self.slot1,
                                     worker_grads, self.hp, self.dev_state, task=self.task,
                                     count=self.count, stream=stream)
            self.global_step += 1          # the authoritative counter is the device one
            return
        if self.count:
            self.ctx.apply(self.optimizer, self.var, self.slot0, self.slot1, worker_grads,
                           self.hp, task=self.task, count=self.count, stream=stream)
        self._finish()

    def apply_gradients_host(self, host_grads, host_var_out=None) -> None:
        """Same step with the gradients in (pinned) host memory and the updated
        variables copied back to host_var_out — the reference-facing call."""
        if self.count:
            self.ctx.ps_step_host(self.optimizer, self.var, self.slot0, self.slot1, host_grads,
                                  self.hp, host_var_out, task=self.task, count=self.count)
        self._finish()

    def apply_async(self, worker_grad, stream=None) -> None:
        """Asynchronous (stale) mode — the example's default (mnist_replica.py:72,
        sync_replicas=False): one worker's push is applied on its own, no aggregation; TF runs
        the optimizer's finish step (Adam's beta powers) with every such apply."""
        self.apply_gradients([worker_grad], stream=stream)

    # ---- checkpoint / restore of this task's shard (TFJobSpec.ModelDir, declared at
    # vendor/github.com/caicloud/kubeflow-clientset/apis/kubeflow/v1alpha1/types.go:46-47 and
    # unused by the reference; the example checkpoints to a temp dir, mnist_replica.py:198-215)
    def save(self, model_dir: str) -> str:
        import os

        import numpy as np
        os.makedirs(model_dir, exist_ok=True)
        if self.dev_state is not None:     # the device holds the authoritative powers and step
            b = self.dev_state[:2].cpu()
            self.hp.beta1_power, self.hp.beta2_power = float(b[0]), float(b[1])
            self.global_step = self.device_global_step()
        path = os.path.join(model_dir, f"ps-{self.task}-of-{self.n_ps}.npz")
        hp = {name: getattr(self.hp, name) for name, _ in self.hp._fields_}
        arrays = {"var": self.var.cpu().numpy(), "global_step": np.int64(self.global_step),
                  "count": np.int64(self.count), "optimizer": np.int64(self.optimizer),
                  "hp_names": np.array(list(hp)), "hp_values": np.array(list(hp.values()),
                                                                        dtype=np.float64)}
        if self.slot0 is not None:
            arrays["slot0"] = self.slot0.cpu().numpy()
        if self.slot1 is not None:
            arrays["slot1"] = self.slot1.cpu().numpy()
        tmp = path + ".tmp.npz"
        np.savez(tmp, **arrays)
        os.replace(tmp, path)       # a reader never sees a half-written checkpoint
        return path

    def restore(self, model_dir: str) -> None:
        import os

        import numpy as np
        path = os.path.join(model_dir, f"ps-{self.task}-of-{self.n_ps}.npz")
        z = np.load(path)
        if int(z["count"]) != self.count or int(z["optimizer"]) != self.optimizer:
            raise ValueError(f"{path} holds a different shard (count {int(z['count'])}, "
                             f"optimizer {int(z['optimizer'])})")
        self.var.copy_(torch.from_numpy(z["var"]))
        if self.slot0 is not None:
            self.slot0.copy_(torch.from_numpy(z["slot0"]))
        if self.slot1 is not None:
            self.slot1.copy_(torch.from_numpy(z["slot1"]))
        import ctypes
        fields = dict(self.hp._fields_)
        for name, value in zip(z["hp_names"], z["hp_values"]):
            is_float = fields[str(name)] is ctypes.c_float
            setattr(self.hp, str(name), float(value) if is_float else int(value))
        self.global_step = int(z["global_step"])
        if self.dev_state is not None:
            self.ctx.opt_state_init(self.dev_state, self.hp, self.global_step)

    def _finish(self) -> None:
        if self.optimizer == accel.OPT_ADAM:
            accel.adam_advance(self.hp)
        self.global_step += 1

    def pull(self) -> torch.Tensor:
        return self.var[:self.count]

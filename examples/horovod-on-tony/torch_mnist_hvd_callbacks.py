"""The Keras-shaped Horovod job (``model.fit`` with callbacks) written against ``tony_b200.parallel.hvd``.

The reference's tony-examples/horovod-on-tony/tensorflow2_keras_mnist.py:55-56,62-83 drives training through
callbacks: ``BroadcastGlobalVariablesCallback(0)`` before the first batch, ``MetricAverageCallback`` after every
epoch, ``LearningRateWarmupCallback`` for the first epochs, and a ``ModelCheckpoint`` appended on rank 0 only.
The same four pieces here, as small callback objects around the fused optimizer:

    python -m tony_b200.cli.cluster_submitter --conf_file examples/horovod-on-tony/tony.xml \
        --executes examples/horovod-on-tony/torch_mnist_hvd_callbacks.py --python_binary_path $(which python)
"""
import argparse
import os
import sys

sys.path.insert(0, os.environ.get("TONY_REPO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")))

import torch

import tony_b200.parallel.hvd as hvd
from tony_b200 import runtime
from tony_b200.models.nn import Linear
from tony_b200.ops import functional as F


class BroadcastGlobalVariables:
    """Every rank starts from root's parameters (ranks are seeded differently on purpose)."""

    def __init__(self, root_rank=0):
        self.root = root_rank

    def on_train_begin(self, model, opt, logs):
        hvd.broadcast_parameters(model.state_dict(), root_rank=self.root)


class MetricAverage:
    """Replace each logged metric by its mean over the ranks (fp32 on the wire)."""

    def on_epoch_end(self, model, opt, logs):
        for k in list(logs):
            logs[k] = hvd.allreduce(torch.tensor([logs[k]], device="cuda"), average=True).item()


class LearningRateWarmup:
    """Ramp from lr to lr * size over the first epochs: a large effective batch from step 0 diverges otherwise."""

    def __init__(self, base_lr, warmup_epochs):
        self.base, self.epochs = base_lr, max(1, warmup_epochs)

    def on_epoch_begin(self, epoch, model, opt, logs):
        frac = min(1.0, (epoch + 1) / self.epochs)
        opt.set_lr(self.base * (1.0 + frac * (hvd.size() - 1)))


class ModelCheckpoint:
    def __init__(self, path):
        self.path = path

    def on_epoch_end(self, model, opt, logs):
        torch.save({k: v.cpu() for k, v in model.state_dict().items()}, self.path)


def fit(model, opt, batches_per_epoch, epochs, make_batch, callbacks):
    logs = {}
    for cb in callbacks:
        getattr(cb, "on_train_begin", lambda *a: None)(model, opt, logs)
    history = []
    for epoch in range(epochs):
        for cb in callbacks:
            getattr(cb, "on_epoch_begin", lambda *a: None)(epoch, model, opt, logs)
        total = 0.0
        for i in range(batches_per_epoch):
            images, labels = make_batch(epoch * batches_per_epoch + i)
            loss = F.softmax_cross_entropy(model(images), labels, vocab=10)
            loss.backward()
            opt.step()
            total += loss.item()
            if epoch == 0 and i == 0:
                runtime.mark_first_step()
        logs = {"loss": total / batches_per_epoch}
        for cb in callbacks:
            getattr(cb, "on_epoch_end", lambda *a: None)(model, opt, logs)
        history.append(logs["loss"])
        if hvd.rank() == 0:
            print(f"epoch {epoch}: mean loss over {hvd.size()} ranks {logs['loss']:.4f}", flush=True)
    return history


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=4)
    ap.add_argument("--batches", type=int, default=10)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--ckpt", default="")
    a = ap.parse_args()
    hvd.init()
    torch.cuda.set_device(hvd.local_rank() % max(torch.cuda.device_count(), 1))
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(99 + hvd.rank())
    model = torch.nn.Sequential()
    model.add_module("fc1", Linear(784, 256, device=dev))
    model.add_module("fc2", Linear(256, 16, device=dev))
    base_lr = 0.02
    opt = hvd.DistributedOptimizer(model, optimizer="sgd", lr=base_lr, momentum=0.9, weight_decay=0.0)
    protos = torch.randn(10, 784, generator=torch.Generator().manual_seed(7))

    def make_batch(step):
        gs = torch.Generator().manual_seed(1000 * step + hvd.rank())
        labels = torch.randint(0, 10, (a.batch,), generator=gs)
        images = (protos[labels] + 0.5 * torch.randn(a.batch, 784, generator=gs)).to(torch.bfloat16)
        return images.to(dev), labels.to(dev)

    callbacks = [BroadcastGlobalVariables(0), MetricAverage(), LearningRateWarmup(base_lr, warmup_epochs=2)]
    if hvd.rank() == 0 and a.ckpt:                  # checkpointing on rank 0 only, like the reference
        callbacks.append(ModelCheckpoint(a.ckpt))
    history = fit(model, opt, a.batches, a.epochs, make_batch, callbacks)
    opt.synchronize()
    opt.check()
    if hvd.rank() == 0:
        assert history[-1] < history[0], history
        print(f"done: loss {history[0]:.4f} -> {history[-1]:.4f}", flush=True)
    if hvd.size() > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()

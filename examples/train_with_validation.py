"""Training shards + a shared validation set on one 8-GPU box, the B200 way.

Two windows per instance (the reference would run two DataPusher processes per consumer,
reference ddl/ddl_env.py:118-124):

  window 1  training images, uint8 CHW, delivered as normalised bf16; every epoch the k-th producers of all
            instances exchange half of their samples (the global shuffle the reference designed,
            reference ddl/shuffle.py:92-108) and shuffle locally with a declared seed (perm_gen on the GPU);
  window 2  a validation set that every consumer must see identically: `replicate=True`.  With
            `exchange="push"` each sample is converted once by the GPU that owns it and stored once through
            the NVLS multicast mapping of the batch rings; NVSwitch writes it into every consumer's ring.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/train_with_validation.py
    python examples/train_with_validation.py        # one GPU, or the CPU execution path without a GPU

Consumer epochs alternate between the windows exactly as in the reference (epoch e reads producer (e mod 2) + 1,
reference ddl/mpi_dataloader.py:213-227), so even epochs train and odd epochs validate.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ddl_b200 as ddl  # noqa: E402

C, H, W = 3, 32, 32
MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
ROWS, BATCH, EPOCHS = 4096, 256, 6


class Shards(ddl.ProducerFunctionSkeleton):
    """Producer 1 of an instance holds training images, producer 2 the validation images."""

    def on_init(self, **kw):
        super().on_init(**kw)
        group = 3                                          # 1 consumer + 2 producers per (virtual) instance
        self.validation = self.rank_global % group == 2
        cv = ddl.Convert.normalize_u8(MEAN, STD, H * W)
        return ddl.DataProducerOnInitReturn(ROWS, C * H * W + 16, (ROWS, C * H * W + 16), (C * H * W, 16),
                                            dtype=np.uint8, convert=cv, shuffle_seed=self.rank_global,
                                            replicate=self.validation)

    def post_init(self, **kw):
        a = kw["my_ary"]
        g = np.random.default_rng(1000 + self.rank_global)
        a[:, : C * H * W] = g.integers(0, 256, size=(ROWS, C * H * W), dtype=np.uint8)
        a[:, C * H * W:] = 0                               # a second field (per-sample metadata) rides in the row


def body(env, conn):
    from ddl_b200.models import TinyConvNet

    loader = ddl.DistributedDataLoader(Shards(), BATCH, conn, EPOCHS, 0.5, "sendrecv_replace", env.color,
                                       env.n_instances)
    dev = env.device
    model = TinyConvNet(10).to(dev)
    if dev.type == "cuda":
        model = model.to(torch.bfloat16)
    if env.n_instances > 1 and dev.type == "cuda":
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[dev.index],
                                                          process_group=env.consumer_group)
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    with loader:
        for epoch in range(EPOCHS):
            validating = epoch % 2 == 1
            seen, loss_sum = 0, 0.0
            for images, extra in loader:                   # device tensors: normalised images + metadata field
                x = images.reshape(BATCH, C, H, W).to(next(model.parameters()).dtype)
                y = (x.float().flatten(1)[:, :8].sum(1).abs() * 7).long() % 10   # synthetic labels from the pixels
                with torch.set_grad_enabled(not validating):
                    loss = torch.nn.functional.cross_entropy(model(x).float(), y)
                if not validating:
                    opt.zero_grad(set_to_none=True)
                    loss.backward()
                    opt.step()
                loss_sum, seen = loss_sum + float(loss.detach()), seen + 1
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
            if env.comm_global.rank == 0:
                print(f"epoch {epoch} ({'validate' if validating else 'train'}): {seen} batches, mean loss {loss_sum / seen:.4f}")


if __name__ == "__main__":
    world = int(os.environ.get("WORLD_SIZE", "1"))
    gpu = torch.cuda.is_available()
    # replicated windows are delivered by the push form; a single rank or the CPU path uses the default pull
    cfg = ddl.LoaderConfig.from_env(exchange="push" if (gpu and world > 1) else "pull", ring_depth=4, prefetch_depth=2,
                                    fetch_streams=2)
    ddl.distributed_dataloader(body, n_instances="world", n_producers=2, config=cfg)()

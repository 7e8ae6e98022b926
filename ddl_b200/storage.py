"""Dedicated storage ranks: GPUs that hold shards and serve them over NVLink, but do not train.

BASELINE config 4 asks for producers:consumers of 1:3 (two producer GPUs feeding six trainers).
The reference cannot express fewer producers than consumers -- every instance needs at least one
producer process of its own (reference ddl/connection.py:29-30).  Here a storage rank plays the
part of the reference's producer processes for the instances assigned to it: it receives their
producer function, runs ``on_init`` / ``post_init`` with the ``rank_global`` the reference would
use, and loads the windows into shards in its HBM.  Because rows never move, that is all it ever
does: consumers' kernels pull the samples they need straight out of its memory.
"""
from __future__ import annotations

import logging

import numpy as np
import torch

from .datapusher import _meta_from_init, _private_copy
from .engine import WindowSpec
from .parallel.heap import SymmetricHeap
from .utils import execute_callbacks

log = logging.getLogger("ddl_b200")


class StorageServer:
    def __init__(self, env, config):
        self.env, self.cfg = env, config
        self.compat = config.compat == "as_wired"

    def serve(self) -> None:
        env = self.env
        comm, me = env.comm_global, env.comm_global.rank
        P, G = env.n_producers, env.n_producers + 1
        hosted = [i for i in range(env.n_instances) if env.storage_rank_of(i) == me]
        work = []
        for i in hosted:
            meta = comm.recv_obj(i, tag=31)
            metas = []
            for k in range(1, P + 1):
                pf = _private_copy(meta.producer_function)
                rank_global = i * G + k
                ret = execute_callbacks("on_init", callbacks=[pf], compat=self.compat, rank_global=rank_global)
                m = _meta_from_init(ret, meta.batch_size)
                metas.append(m)
                work.append((i, k - 1, pf, rank_global, WindowSpec(m)))
            comm.send_obj(metas, i, tag=32)

        heap = SymmetricHeap(comm, env.device)
        cuda = torch.device(env.device).type == "cuda"
        shards = [heap.alloc(f"shard{k}_i{i}", s.nData * s.row_bytes, zero=False) for i, k, _, _, s in work]
        heap.publish()
        for (i, k, pf, rank_global, s), shard in zip(work, shards):
            if cuda:
                buf, _ = heap.alloc_pinned(f"window{k}_i{i}", s.nData * s.window_bytes)
                window = buf.numpy(s.np_dtype, (s.nData, s.window_values))
            else:
                window = np.zeros((s.nData, s.window_values), dtype=s.np_dtype)
            execute_callbacks("post_init", callbacks=[pf], compat=self.compat, rank_global=rank_global,
                              my_ary=window, n_instance=env.n_instances)
            dst = shard.tensor(torch.uint8, (s.nData, s.row_bytes))
            dst[:, :s.window_bytes].copy_(torch.from_numpy(window.reshape(s.nData, -1).view(np.uint8)))
            if s.var:  # variable-size samples: the consumer needs the valid length of every row
                lens = getattr(pf, "lengths", None)
                comm.send_obj(None if lens is None else np.asarray(lens).astype(np.int32), i, tag=40 + k)
        if cuda:
            torch.cuda.synchronize(env.device)
        log.info("[%03d] storage rank holds %d shards for instances %s", me, len(work), hosted)
        comm.barrier()  # shards resident
        comm.barrier()  # every consumer has finished
        heap.close()

"""The global shuffle: partner selection and the cross-instance exchange
(reference ddl/shuffle.py:5-108).

Every producer iteration the k-th producers of all instances pick a random permutation cycle
structure -- a derangement without 2-cycles -- and swap the leading ``num_exchange`` rows of
their windows with two partners: rows ``[0, ne/2)`` come from the rank that sends to me, rows
``[ne/2, ne)`` from the rank I send to.

* ``SendRecvReplaceGlobalShuffler`` keeps the reference's name and constructor and moves
  *rows* of a host numpy window through the control-plane communicator: the like-for-like
  replacement of its two ``Sendrecv_replace`` calls (process transport, host windows).
* ``DeviceGlobalShuffler`` applies the same exchange to slot tables in HBM with the
  ``index_exchange`` kernel: 8 bytes per sample cross NVLink instead of the sample, which
  travels once, later, inside ``exchange_gather_collate``.

Partner choice is bit-exact with the reference (native ``bddl_partner_select``; test vectors in
SURVEY.md 3.5).
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from .ops.rng import Pcg64


class GlobalShuffler:
    """Partner selection shared by every exchange method (reference ddl/shuffle.py:5-79)."""

    def __init__(self, my_ary: Optional[np.ndarray], num_exchange: int, comm_nth_pusher, comm_global,
                 seed: Optional[int] = None, n_instance: Optional[int] = None, my_index: Optional[int] = None):
        self.my_ary = my_ary
        self.num_exchange = int(num_exchange)
        self.comm_nth_pusher = comm_nth_pusher
        self.comm_global = comm_global
        self.rank = comm_global.Get_rank() if comm_global is not None else 0
        self.n_instance = int(n_instance if n_instance is not None else comm_nth_pusher.Get_size())
        self.my_index = int(my_index if my_index is not None else comm_nth_pusher.Get_rank())
        if seed is None:
            # position inside the instance group: identical on all members of comm_nth_pusher
            seed = self.rank % (comm_global.Get_size() // self.n_instance)
        self.seed = int(seed)
        self.comm_shuffle_rng = Pcg64(self.seed)
        self.iterations = 0
        self.rejected_draws = 0
        self.last_permutation: Optional[np.ndarray] = None

    def _calculate_comm_partner(self) -> Tuple[int, int]:
        """(send_to, recv_from) for this iteration; advances the partner RNG."""
        send_to, recv_from, perm, tries = self.comm_shuffle_rng.partner(self.n_instance, self.my_index)
        self.iterations += 1
        self.rejected_draws += tries
        self.last_permutation = perm
        return send_to, recv_from

    # -- checkpoint ---------------------------------------------------------------------------------
    def state_dict(self) -> dict:
        return {"rng": self.comm_shuffle_rng.state_dict(), "iterations": self.iterations,
                "rejected_draws": self.rejected_draws}

    def load_state_dict(self, d: dict) -> None:
        self.comm_shuffle_rng.load_state_dict(d["rng"])
        self.iterations = d.get("iterations", 0)
        self.rejected_draws = d.get("rejected_draws", 0)


class SendRecvReplaceGlobalShuffler(GlobalShuffler):
    """Row exchange on a host window (reference ddl/shuffle.py:82-108)."""

    def global_shuffle(self, **kwargs) -> None:
        send_to, recv_from = self._calculate_comm_partner()
        if self.n_instance == 1 or self.num_exchange == 0:
            return
        half = self.num_exchange // 2
        self.comm_nth_pusher.sendrecv_replace(self.my_ary[:half], dest=send_to, source=recv_from,
                                              sendtag=50, recvtag=50)
        self.comm_nth_pusher.sendrecv_replace(self.my_ary[half:self.num_exchange], dest=recv_from,
                                              source=send_to, sendtag=51, recvtag=51)


class DeviceGlobalShuffler(GlobalShuffler):
    """The same exchange on slot tables; ``engine`` owns the tables and issues the kernel."""

    def __init__(self, engine, producer_idx: int, num_exchange: int, n_instance: int, my_index: int, seed: int):
        super().__init__(None, num_exchange, None, None, seed=seed, n_instance=n_instance, my_index=my_index)
        self.engine = engine
        self.producer_idx = producer_idx

    def global_shuffle(self, **kwargs) -> None:
        send_to, recv_from = self._calculate_comm_partner()
        self.engine.exchange_tables(self.producer_idx, send_to, recv_from, self.num_exchange)


def make_global_shuffler(method: str, *args, **kwargs) -> GlobalShuffler:
    """The reference accepts exactly one method string (reference ddl/datapusher.py:96-106)."""
    if method == "sendrecv_replace":
        return SendRecvReplaceGlobalShuffler(*args, **kwargs)
    raise NotImplementedError(f"{method} is not implemented")

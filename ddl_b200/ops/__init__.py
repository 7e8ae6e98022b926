"""Device ops of the loader: RNG, slot-table shuffles and the fused batch-assembly kernel."""
from .checksum import BatchChecksum, checksum_reference
from .rng import Pcg64
from .gather import Convert, GatherPlan, SourceTable, gather_reference, make_slot, split_slot, round_up
from .tables import index_exchange, perm_gen, rng_states_from_tensor, rng_states_to_tensor, slot_iota

__all__ = [
    "Pcg64", "BatchChecksum", "checksum_reference", "Convert", "GatherPlan", "SourceTable", "gather_reference", "make_slot", "split_slot", "round_up",
    "index_exchange", "perm_gen", "rng_states_from_tensor", "rng_states_to_tensor", "slot_iota",
]

"""tatajuba_b200 — B200-native (sm_100a) replacement for tatajuba's hot path (HT scan -> context histograms)."""
from .hopo_counter import (Context, HopoCounter, TatajubaOptions, TjbError, finalise_hopo_counter,  # noqa: F401
                           new_or_append_hopo_counter_from_file, HOPO_DTYPE)

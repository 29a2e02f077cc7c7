"""ctypes binding of libgeobpe_b200.so (C ABI in include/geobpe_b200.h).

The product path has no CPU fallback: a missing library or a missing CUDA device raises."""
import ctypes as C
import os

import torch

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libgeobpe_b200.so")

P, I64, I32, U32, U64, F64 = C.c_void_p, C.c_int64, C.c_int, C.c_uint32, C.c_uint64, C.c_double

# name -> argtypes (all return int unless listed in _RESTYPES)
SIGNATURES = {
    "gb_residue_backbone": [P, I64, I64, P, P],
    "gb_select_flagged": [P, I64, U32, P, P, P, P],
    "gb_select_eq_u64": [P, I64, U64, P, P, P, P],
    "gb_select_flagged_static": [P, I64, U32, P, P, P],
    "gb_select_flagged_static_max_n": [],
    "gb_col_minmax": [P, I64, I64, P, P, P],
    "gb_col_hist": [P, I64, I64, P, I32, P, P, P],
    "gb_quantize_residues": [P, I64, I64, P, P, I64, P, I32, P, I32, F64, F64, F64, P, P, P, P, P],
    "gb_table_clear": [P, P, P, I64, P],
    "gb_pair_count_m": [P, P, I64, P, P, P, I64, P, P],
    "gb_pair_count_h": [P, P, P, I64, P, U64, P, P, P, I64, P, P],
    "gb_first_positions_m": [P, P, I64, P, P, P, I64, P, P, P, I32, P],
    "gb_first_positions_h": [P, P, P, I64, P, U64, P, P, P, I64, P, P, P, I32, P],
    "gb_key_count": [P, I64, P, P, P, I64, P, P],
    "gb_table_argmax": [P, P, P, I64, P, P, P, I32, P],
    "gb_table_lookup": [P, P, P, I64, P, I32, P, P, P],
    "gb_table_dump": [P, P, P, I64, P, P, P, P, P],
    "gb_table_to_slots": [P, P, I64, P, I64, P],
    "gb_table_dump_marked": [P, P, I64, P, I64, P, P, P, I32, P],
    "gb_slot_count_m": [P, P, I64, P, I32, P, I32, I32, P],
    "gb_slot_count_h": [P, P, P, I64, P, U64, P, I32, P, I32, I32, P],
    "gb_slots_top": [P, I32, U32, P, P, I32, P, P],
    "gb_cand_count_m": [P, P, I64, I32, P, P, I32, P, P, P, I64, P, I32, P],
    "gb_cand_count_h": [P, P, P, I64, P, U64, I32, P, P, I32, P, P, P, I64, P, I32, P],
    "gb_flag_pairs_m": [P, P, I64, U64, P, P],
    "gb_flag_pairs_h": [P, P, P, I64, P, U64, U64, P, P],
    "gb_resolve_overlaps": [P, I64, P],
    "gb_scatter_u32": [P, P, I64, P, P],
    "gb_merge_compact": [P, I64] + [P] * 6 + [P] * 6 + [P, P, P, U64, U32, P, P, P],
    "gb_write_events": [P, I64, P, P, P, P, P, P, P],
    "gb_verify_flagged_h": [P, I64, P, P, P, I64, U32, U32, P, P],
    "gb_pack_result": [P, P, P, P, P, P, I32, P, P, P],
    "gb_oplog_m": [P, I64, P, P, P, P, I64, P, P, P, P, P],
    "gb_pair_keys_m": [P, P, P, I64, P, P, P],
    "gb_verify_windows_h": [P, I64, P, P, P, I64, U32, U32, P, P],
    "gb_window_coords": [P, I64, I64, P, P, P, I64, I32, I32, P, P, F64, F64, F64, P, P, I32, P, P],
    "gb_kabsch_assign": [P, I64, I32, P, I32, P, P, P],
    "gb_rmsd_matrix": [P, I32, I32, P, P],
    "gb_kmedoids_pam": [P, I32, I32, P, I32, P, P, P],
    "gb_pam_max_n": [],
    "gb_pam_max_k": [],
    "gb_pairwise_sum_host": [P, I32],
    "gb_h_loop": [P, I32, I32, I32, I32, I32, P, P],
    "gb_h_loop_nargs": [],
    "gb_h_loop2": [P, I32, I32, I32, P],
    "gb_h_loop2_nargs": [],
    "gb_induce_replay": [P] * 9 + [I64, P, I64, P, P, P, I32] + [P] * 6 + [P, P, P] + [P] * 5 + [I64, P, P, P, P],
    "gb_induce_max_atoms": [],
    "gb_reconstruction": [P, I64, P, I32, P, P, P, I64, P, I32, P, I64, P, P, I64, F64, F64, F64, P, P, P, P, P, P, P],
    "gb_lddt_ca": [P, P, P, I64, F64, P, P],
    "gb_token_stream": [P, P, P, P, I64, U32, U32, U32, U32, P, P],
    "gb_rehash_tokens": [P, P, P, P, I64, U64, P, P],
    "gb_h_loop2_state_words": [],
    "gb_h_loop2_delta_bytes": [I64, I32],
    "gb_h_loop2_export_bytes": [I64, I32],
    "gb_h_loop2_flag_bytes": [],
    "gb_h_loop2_max_world": [],
    "gb_h_loop_part": [P, I32, I32, I32, P],
    "gb_h_loop_xbuf_elems": [],
    "gb_h_loop_sbuf_bytes": [I32],
    "gb_peer_alloc": [I64, P, P],
    "gb_peer_open": [P, P],
    "gb_peer_close": [P],
    "gb_peer_free": [P],
    "gb_peer_max_world": [],
    "gb_h_loop_use_graphs": [I32],
    "gb_h_loop_prepare": [P, I32, P],
    "gb_h_loop_kernels_per_step": [],
    "gb_h_loop_time_k1": [I32],
    "gb_h_loop_incremental": [I32],
    "gb_h_loop_round2": [I32],
    "gb_h_loop_k1_times": [P, I32],
    "gb_last_error": [],
    "gb_abi_version": [],
    "gb_sm_count": [],
    "gb_select_scratch_elems": [I64],
}
_RESTYPES = {"gb_last_error": C.c_char_p, "gb_pairwise_sum_host": C.c_float, "gb_select_scratch_elems": I64, "gb_select_flagged_static_max_n": I64, "gb_h_loop2_delta_bytes": I64,
             "gb_h_loop2_export_bytes": I64, "gb_h_loop2_flag_bytes": I64, "gb_h_loop_xbuf_elems": I64, "gb_h_loop_sbuf_bytes": I64}

_lib = None
launches = 0  # kernels launched through the C ABI (bench.py reports it)
# kernels per entry point (default 1)
KERNELS = {"gb_select_flagged": 3, "gb_select_flagged_static": 3, "gb_select_eq_u64": 3, "gb_merge_compact": 3, "gb_table_argmax": 2,
           "gb_col_minmax": 3, "gb_first_positions_m": 2, "gb_first_positions_h": 2}


def load():
    """dlopen the in-tree library and declare every prototype; raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m geobpe_b200.build` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.argtypes = args
        fn.restype = _RESTYPES.get(name, C.c_int)
    _lib = lib
    return lib


def ptr(t):
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "C ABI takes contiguous CUDA tensors"
    return t.data_ptr()


def stream():
    """raw handle of torch's current stream (the C accessor: torch.cuda.current_stream() costs ~17 us per call)"""
    return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device())


def call(name, *args):
    """Invoke a status-returning entry point on torch's current stream; raise RuntimeError on failure."""
    global launches
    lib = load()
    rc = getattr(lib, name)(*args, stream())
    launches += KERNELS.get(name, 1)
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.gb_last_error().decode()}")


def call_n(name, nkernels, *args):
    """call() for an entry point that enqueues `nkernels` kernels"""
    global launches
    lib = load()
    rc = getattr(lib, name)(*args, stream())
    launches += nkernels
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.gb_last_error().decode()}")


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("geobpe_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    load()

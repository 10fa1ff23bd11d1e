"""ctypes binding of include/lime_cuda.h (liblime_b200.so).

The shared library is the product: it is built in-tree by ``__graft_entry__.build()`` /
``make -C lime_b200/csrc`` and loaded from ``lime_b200/liblime_b200.so``.  There is no
Python or CPU fallback -- if the library is missing or no CUDA device is present the
import of a context fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# LIME_B200_LIB: another build of the same library (kernel tuning experiments, tools/gpu_k12_variants.sh)
LIB_PATH = os.environ.get("LIME_B200_LIB") or os.path.join(_HERE, "liblime_b200.so")

OK = 0
ERR_ARG, ERR_CUDA, ERR_MAXLEN, ERR_HALO, ERR_STATE, ERR_NOMEM, ERR_ROWS, ERR_RETRY = -1, -2, -3, -4, -5, -6, -7, -8
MEM_HOST, MEM_DEVICE = 0, 1


class LimeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"lime error {code}: {msg}")
        self.code = code


u8p, u32p, u64p, f32p = (C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64),
                         C.POINTER(C.c_float))
vp = C.c_void_p

# name -> (restype, argtypes); must list every symbol include/lime_cuda.h declares
IPC_HANDLE_BYTES = 64       # LIME_IPC_HANDLE_BYTES
SIGNATURES = {
    "lime_init": (C.c_int, [C.POINTER(vp), C.c_int]),
    "lime_destroy": (None, [vp]),
    "lime_last_error": (C.c_char_p, [vp]),
    "lime_version": (C.c_char_p, []),
    "lime_set_stream": (C.c_int, [vp, vp]),
    "lime_pinned_alloc": (vp, [C.c_uint64]),
    "lime_pinned_free": (None, [vp]),
    "lime_load": (C.c_int, [vp, vp, vp, vp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int,
                            C.c_uint32, C.c_uint32, C.c_int]),
    "lime_set_alpha_limit": (C.c_int, [vp, C.c_uint32]),
    "lime_set_alpha": (C.c_int, [vp, C.c_uint32]),
    "lime_host_pack_lcp": (C.c_int, [vp, vp, C.c_uint64, C.c_uint32, C.c_int]),
    "lime_host_pack_da3": (C.c_int, [vp, vp, C.c_uint64]),
    "lime_host_pack_lcp_da3_pipelined": (C.c_int, [vp, vp, vp, vp, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int, C.c_int, C.c_int]),
    "lime_host_pack_lcp_pipelined": (C.c_int, [vp, vp, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int, C.c_int, C.c_int]),
    "lime_host_threads": (C.c_int, []),
    "lime_host_narrow_lcp": (C.c_int, [vp, vp, C.c_uint64, C.c_int]),
    "lime_host_narrow_lcp_pipelined": (C.c_int, [vp, vp, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int]),
    "lime_cluster_detect": (C.c_int, [vp, C.c_uint32, u64p, u64p]),
    "lime_clusters_copy": (C.c_int, [vp, vp, C.c_uint64]),
    "lime_clusters_set": (C.c_int, [vp, vp, C.c_uint64, C.c_uint64]),
    "lime_similarity": (C.c_int, [vp, C.c_int]),
    "lime_cluster_similarity": (C.c_int, [vp, C.c_uint32, C.c_int, C.c_int, u64p, u64p]),
    "lime_sim_bind": (C.c_int, [vp, vp, C.c_uint64, C.c_uint64]),
    "lime_sim_pitch": (C.c_uint64, [C.c_uint32]),
    "lime_sim_copy": (C.c_int, [vp, vp]),
    "lime_sim_export": (C.c_int, [vp, C.c_uint32, C.c_uint32, vp, C.c_uint64, u64p]),
    "lime_sim_import": (C.c_int, [vp, vp, C.c_uint64]),
    "lime_choose": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_float, C.c_uint32, C.c_uint32, u64p]),
    "lime_choose_copy": (C.c_int, [vp, C.c_int, vp, vp]),
    "lime_slot_set": (C.c_int, [vp, C.c_int, vp, C.c_uint64, vp, C.c_uint32]),
    "lime_classify": (C.c_int, [vp, C.c_int, C.c_uint32, vp, vp, C.c_int, C.c_int, vp, vp, vp]),
    "lime_classify_result": (C.c_int, [vp, vp, vp, vp, C.c_int]),
    "lime_xchg_alloc": (C.c_int, [vp, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_uint64, C.c_uint64, vp]),
    "lime_xchg_open": (C.c_int, [vp, vp]),
    "lime_xchg_close": (C.c_int, [vp]),
    "lime_xchg_accumulate": (C.c_int, [vp, C.c_uint32, C.c_int, C.c_int]),
    "lime_xchg_apply": (C.c_int, [vp, C.c_int]),
    "lime_xchg_choose": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_float, C.c_uint32]),
    "lime_xchg_status": (C.c_int, [vp, C.c_int, u64p]),
    "lime_choose_static": (C.c_int, [vp, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, vp, C.c_uint64, vp, vp, vp, vp,
                                     C.c_char_p, C.c_uint64]),
    "lime_strand_run": (C.c_int, [vp, C.c_int, C.c_uint32, C.c_int, C.c_uint32, C.c_float]),
    "lime_strand_status": (C.c_int, [vp, C.c_int, u64p]),
    "lime_launch_count": (C.c_uint64, [vp]),
    "lime_prof_enable": (C.c_int, [vp, C.c_int]),
    "lime_prof_kernels": (C.c_int, []),
    "lime_prof_get": (C.c_int, [vp, C.c_int, C.POINTER(C.c_char_p), u64p, C.POINTER(C.c_double)]),
    "lime_prof_reset": (C.c_int, [vp]),
    "lime_sim_stats": (C.c_int, [vp, u64p]),
    # host-side helpers (lime_host.h)
    "lime_host_fix_rank": (C.c_int64, [C.c_char_p, C.c_int, C.c_uint32, vp, C.c_uint64, vp]),
    "lime_host_write_csv": (C.c_int, [C.c_char_p, C.c_uint32, vp, vp, vp]),
    "lime_host_read_gesa": (C.c_int64, [C.c_char_p, vp, vp, vp, C.c_uint64]),
    "lime_host_write_res_txt": (C.c_int, [C.c_char_p, vp, C.c_uint64, vp, C.c_uint32]),
    "lime_host_read_res_txt": (C.c_int64, [C.c_char_p, C.c_uint32, vp, C.c_uint64, vp]),
}

_lib = None


def load():
    """Load liblime_b200.so and attach prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C lime_b200/csrc` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib

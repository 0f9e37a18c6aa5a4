"""ctypes binding of libsevenn_b200.so (the C ABI in include/sevenn_b200.h).

There is no fallback: if the library is missing, or a CUDA device is not
available, the product raises.  (tests/ may inject an emulation build of the
same sources to check host logic on a box without a GPU; the package itself
never loads it.)
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, '_build', 'libsevenn_b200.so')

c_p, c_i, c_ll, c_f = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


class GemmArgs(ctypes.Structure):
    _fields_ = [('A', c_p), ('sAn', c_ll), ('sAm', c_i), ('B', c_p), ('ldb', c_i), ('C', c_p), ('sCn', c_ll),
                ('sCm', c_i), ('M', c_ll), ('R', c_i), ('N', c_i), ('K', c_i), ('epi', c_i), ('act', c_i),
                ('cst', c_f), ('C2', c_p), ('P', c_p)]


class GatePiece(ctypes.Structure):
    _fields_ = [('kind', c_i), ('l', c_i), ('mul', c_i), ('in_off', c_i), ('in_stride', c_i), ('gate_off', c_i),
                ('out_off', c_i), ('out_stride', c_i), ('act', c_i), ('cst', c_f)]


SIGNATURES = {
    'sevenn_abi_version': ([], c_i),
    'sevenn_scan_i32': ([c_p, c_p, c_i, c_p, c_p], c_i),
    'sevenn_scan_tmp_ints': ([c_i], c_ll),
    'sevenn_nl_bin': ([c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p], c_i),
    'sevenn_nl_scatter': ([c_p, c_i, c_p, c_p, c_p, c_p], c_i),
    'sevenn_nl_count': ([c_p, c_p, c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_p], c_i),
    'sevenn_nl_fill': ([c_p, c_p, c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p], c_i),
    'sevenn_csc_build': ([c_p, c_ll, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p], c_i),
    'sevenn_edge_fwd': ([c_p, c_ll, c_i, c_i, c_i, c_i, c_f, c_f, c_f, c_p, c_p, c_p, c_p], c_i),
    'sevenn_edge_bwd': ([c_p, c_ll, c_i, c_i, c_i, c_i, c_f, c_f, c_f, c_p, c_p, c_i, c_ll, c_p, c_p, c_p], c_i),
    'sevenn_edge_fwd_tab': ([c_p, c_ll, c_i, c_i, c_f, c_i, c_p, c_p], c_i),
    'sevenn_edge_bwd_tab': ([c_p, c_ll, c_i, c_i, c_f, c_p, c_i, c_ll, c_p, c_p], c_i),
    'sevenn_gemm_f32': ([ctypes.POINTER(GemmArgs), c_p], c_i),
    'sevenn_gemm_f32_grouped': ([c_i, ctypes.POINTER(GemmArgs), c_p], c_i),
    'sevenn_gemm_tc_plan': ([c_i, c_i, ctypes.POINTER(c_i), ctypes.POINTER(c_i), ctypes.POINTER(c_i),
                             ctypes.POINTER(c_ll)], c_i),
    'sevenn_gemm_tc': ([ctypes.POINTER(GemmArgs), c_p, c_p], c_i),
    'sevenn_gemm_tc_image': ([c_p, c_i, c_i, c_p, c_p], c_i),
    'sevenn_gemm_tc_grouped': ([c_i, ctypes.POINTER(GemmArgs), ctypes.POINTER(c_p), c_p], c_i),
    'sevenn_linear_by_type': ([c_p, c_ll, c_i, c_p, c_p, c_p, c_ll, c_i, c_ll, c_i, c_i, c_i, c_i, c_p], c_i),
    'sevenn_embed': ([c_p, c_ll, c_i, c_p, c_p, c_p], c_i),
    'sevenn_gate_fwd': ([c_p, c_i, c_p, c_i, c_ll, ctypes.POINTER(GatePiece), c_i, c_p], c_i),
    'sevenn_gate_bwd': ([c_p, c_i, c_p, c_i, c_p, c_ll, ctypes.POINTER(GatePiece), c_i, c_p], c_i),
    'sevenn_rescale': ([c_p, c_p, c_ll, c_p, c_p, c_i, c_p, c_p, c_p], c_i),
    'sevenn_colsum_f64': ([c_p, c_ll, c_i, c_p, c_p, c_p], c_i),
    'sevenn_colsum_f64_noalloc': ([c_p, c_ll, c_i, c_p, c_p], c_i),
    'sevenn_edge_fwd_tab_dev': ([c_p, c_ll, c_p, c_i, c_i, c_f, c_i, c_p, c_p], c_i),
    'sevenn_edge_bwd_tab_dev': ([c_p, c_ll, c_p, c_i, c_i, c_f, c_p, c_i, c_ll, c_p, c_p], c_i),
    'sevenn_csc_build_dev': ([c_p, c_ll, c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p], c_i),
    'sevenn_nl_fill_capped': ([c_p, c_p, c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_ll, c_p, c_p, c_p, c_p, c_p, c_p, c_p,
                               c_p], c_i),
    'sevenn_model_create': ([c_p, c_ll, c_p, c_ll, c_p], c_i),
    'sevenn_model_destroy': ([c_p], c_i),
    'sevenn_model_set_tc_min_rows': ([c_p, c_i], c_i),
    'sevenn_model_launches': ([c_p], c_ll),
    'sevenn_model_tc_launches': ([c_p], c_ll),
    'sevenn_model_workspace_bytes': ([c_p, c_ll, c_ll, c_ll, c_i], c_ll),
    'sevenn_model_count_edges': ([c_p, c_ll, c_ll, c_p, c_p, c_p, c_p, c_p, c_ll, c_p, c_p], c_i),
    'sevenn_model_edges_workspace_bytes': ([c_p, c_ll, c_ll], c_ll),
    'sevenn_model_evaluate_edges': ([c_p, c_ll, c_p, c_p, c_p, c_ll, c_p, c_p, c_p, c_ll, c_p, c_p], c_i),
    'sevenn_model_evaluate_graph': ([c_p, c_p, c_p, c_p, c_ll, c_p, c_p], c_i),
    'sevenn_model_evaluate_positions': ([c_p, c_ll, c_ll, c_p, c_p, c_p, c_p, c_p, c_ll, c_p, c_p, c_ll, c_p, c_p], c_i),
    'sevenn_act_mul': ([c_p, c_p, c_ll, c_i, c_f, c_i, c_p, c_p], c_i),
    'sevenn_gate_bwd2': ([c_p, c_i, c_p, c_i, c_p, c_ll, ctypes.POINTER(GatePiece), c_i, c_p, c_p, c_p], c_i),
    'sevenn_edge_train': ([c_p, c_ll, c_i, c_i, c_i, c_i, c_f, c_f, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_p], c_i),
    'sevenn_gemm_tn_splits': ([c_ll], c_i),
    'sevenn_gemm_tn': ([c_p, c_ll, c_i, c_p, c_ll, c_i, c_ll, c_i, c_i, c_i, c_p, c_i, c_f, c_i, c_p, c_p, c_p], c_i),
    'sevenn_adam_step': ([c_p, c_p, c_p, c_p, c_ll, c_i, c_f, c_f, c_f, c_f, c_f, c_p], c_i),
    'sevenn_adam_step_dev': ([c_p, c_p, c_p, c_p, c_ll, c_p, c_f, c_f, c_f, c_f, c_f, c_p], c_i),
    'sevenn_force_virial': ([c_p, c_p, c_p, c_i, c_p, c_p, c_ll, c_p, c_p, c_p], c_i),
    'sevenn_segment_gather_sum': ([c_p, c_i, c_p, c_p, c_ll, c_p, c_i, c_p], c_i),
    'sevenn_gather_rows': ([c_p, c_i, c_p, c_ll, c_p, c_p], c_i),
    'sevenn_scatter_add_rows': ([c_p, c_i, c_p, c_ll, c_p, c_p], c_i),
    'sevenn_push_rows_peers': ([c_p, c_i, c_p, c_ll, c_ll, c_i, c_p, c_p, c_p, c_p], c_i),
}
CONV_FWD_ARGS = [c_p, c_i, c_p, c_p, c_p, c_p, c_i, c_f, c_p, c_i, c_p]
CONV_BWD_ARGS = [c_p, c_i, c_p, c_p, c_p, c_p, c_i, c_f, c_p, c_i, c_p, c_p, c_i, c_p, c_ll, c_p]
CONV_BWD_TAB_ARGS = [c_p, c_i, c_p, c_p, c_p, c_p, c_i, c_f, c_p, c_i, c_p, c_i, c_p, c_ll, c_p]


class SevennError(RuntimeError):
    pass


class Lib:
    """Typed view of the shared library; every call checks the status code."""

    def __init__(self, cdll):
        self.cdll = cdll
        for name, (args, res) in SIGNATURES.items():
            if hasattr(cdll, name):
                f = getattr(cdll, name)
                f.argtypes, f.restype = args, res
        self._conv = {}

    def call(self, name, *args):
        rc = getattr(self.cdll, name)(*args)
        if rc != 0:
            raise SevennError(f'{name} failed with status {rc}')

    def fused_fwd(self, sig):
        """Fused radial-MLP-last-layer + convolution forward of this signature, or None."""
        key = ('fused', sig)
        if key not in self._conv:
            f = getattr(self.cdll, f'sevenn_conv_fwd_fused_{sig}', None) if hasattr(
                self.cdll, f'sevenn_conv_fwd_fused_{sig}') else None
            if f is not None:
                f.argtypes = [c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_i, c_i, c_f, c_p, c_i, c_p, c_p]
                f.restype = c_i
            self._conv[key] = f
        return self._conv[key]

    def conv_tab(self, sig):
        """Tabulated-radial forward / backward of this signature (same translation unit as conv())."""
        key = ('tab', sig)
        if key not in self._conv:
            f = getattr(self.cdll, f'sevenn_conv_fwd_tab_{sig}')
            b = getattr(self.cdll, f'sevenn_conv_bwd_tab_{sig}')
            f.argtypes, f.restype = CONV_FWD_ARGS, c_i
            b.argtypes, b.restype = CONV_BWD_TAB_ARGS, c_i
            self._conv[key] = (f, b)
        return self._conv[key]

    def has_conv(self, sig):
        return hasattr(self.cdll, f'sevenn_conv_fwd_{sig}')

    def conv(self, sig):
        if sig not in self._conv:
            f = getattr(self.cdll, f'sevenn_conv_fwd_{sig}')
            b = getattr(self.cdll, f'sevenn_conv_bwd_{sig}')
            f.argtypes, f.restype = CONV_FWD_ARGS, c_i
            b.argtypes, b.restype = CONV_BWD_ARGS, c_i
            self._conv[sig] = (f, b)
        return self._conv[sig]


_cached = None


def load() -> Lib:
    """The CUDA library.  Raises if it has not been built or no GPU is present."""
    global _cached
    if _cached is None:
        import torch
        if not os.path.exists(LIB_PATH):
            raise SevennError(f'{LIB_PATH} is missing: run `python -m sevennet_exp_b200.build` '
                              '(or __graft_entry__.build()); there is no CPU fallback')
        if not torch.cuda.is_available():
            raise SevennError('sevennet_exp_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback')
        _cached = Lib(ctypes.CDLL(LIB_PATH))
    return _cached


def load_for_symbol_check() -> ctypes.CDLL:
    """dlopen without requiring a GPU (tests: the ABI exports what the header declares)."""
    if not os.path.exists(LIB_PATH):
        raise SevennError(f'{LIB_PATH} is missing')
    return ctypes.CDLL(LIB_PATH)

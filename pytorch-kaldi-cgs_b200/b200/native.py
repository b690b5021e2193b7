"""ctypes binding of ``libhcgs_b200.so`` (the C ABI declared in ``include/hcgs_b200.h``).

This is the only place Python touches the native library.  There is deliberately no
fallback: if the shared object is missing, or a compute entry point is called without
a CUDA device, an exception is raised -- the product path never silently runs on the CPU.
PyTorch is used for device memory and streams only (``tensor.data_ptr()``,
``torch.cuda.current_stream()``).
"""
import collections
import ctypes
import hashlib
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int64, c_void_p

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), 'libhcgs_b200.so')

ACT_CODES = {'linear': 0, 'relu': 1, 'tanh': 2, 'htanh': 3, 'sigmoid': 4, 'leaky_relu': 5, 'elu': 6}
MAX_GROUPS = 4


class NativeError(RuntimeError):
    pass


_lib = None

_f = c_void_p          # device pointer to float (passed as integer address)
_pp = POINTER(c_void_p)  # host array of device pointers

# name -> (restype, argtypes); kept in one table so tests can check it against the header
SIGNATURES = {
    'hcgs_abi_version': (c_int, []),
    'hcgs_last_error': (c_char_p, []),
    'hcgs_device_info': (c_int, [c_int, POINTER(c_int)]),
    'hcgs_plan_create': (c_int, [POINTER(c_float), c_int, c_int, c_int, c_int, POINTER(c_void_p)]),
    'hcgs_plan_destroy': (c_int, [c_void_p]),
    'hcgs_plan_info': (c_int, [c_void_p, POINTER(c_int64)]),
    'hcgs_mask_apply': (c_int, [_pp, c_int, _f, c_int64, c_void_p]),
    'hcgs_quantize_w': (c_int, [_f, _f, c_int64, c_int, c_void_p]),
    'hcgs_quantize_act': (c_int, [_f, c_int, _f, c_int, c_int64, c_int, c_int, c_int, _f, c_void_p]),
    'hcgs_bsgemm_nt': (c_int, [c_void_p, c_int, _pp, c_int, _f, c_int, c_int64, _pp, _f, c_int, c_void_p]),
    'hcgs_bsgemm_nn': (c_int, [c_void_p, c_int, _pp, c_int, _f, c_int, c_int64, _f, c_int, c_void_p]),
    'hcgs_bs_sddmm': (c_int, [c_void_p, c_int, _f, c_int, _f, c_int, c_int64, _pp, c_int, _f, c_void_p]),
    'hcgs_colsum': (c_int, [_f, c_int, c_int64, c_int, c_int, _pp, c_void_p]),
    'hcgs_bn_fwd': (c_int, [_f, c_int, _f, c_int, c_int64, c_int, c_int, _pp, _pp, _pp, _pp, c_float, c_float,
                            c_int, _f, _f, _f, c_int64, c_void_p]),
    'hcgs_bn_fwd_static': (c_int, [_f, c_int, _f, c_int, c_int64, c_int, c_int, _pp, _pp, _pp, _pp, c_float, c_float,
                                   c_int, _f, _f, c_void_p]),
    'hcgs_bn_bwd': (c_int, [_f, c_int, _f, c_int, c_int64, c_int, c_int, _pp, _f, _f, _f, c_int, _pp, _pp, _f,
                            c_int64, c_void_p]),
    'hcgs_bn_fwd_il': (c_int, [_f, _f, c_int64, c_int, c_int, c_int, _pp, _pp, _pp, _pp, c_float, c_float, c_int, _f,
                               _f, _f, c_int64, c_void_p]),
    'hcgs_bn_bwd_il': (c_int, [_f, _f, _f, c_int64, c_int, c_int, c_int, _pp, _f, _f, _f, _f, _pp, _pp, _f, c_int64,
                               c_void_p]),
    'hcgs_bn_stats_il': (c_int, [_f, c_int64, c_int64, c_int64, c_int, c_int, c_int, _pp, _pp, c_float, c_float, c_int,
                                 _f, _f, c_void_p]),
    'hcgs_layernorm_fwd': (c_int, [_f, c_int64, _f, c_int64, c_int64, c_int, _f, _f, c_float, _f, _f, c_void_p]),
    'hcgs_layernorm_bwd': (c_int, [_f, c_int64, _f, c_int64, c_int64, c_int, _f, c_float, _f, _f, _f, c_int64, _f, _f,
                                   c_void_p]),
    'hcgs_lstm_pack': (c_int, [c_void_p, _pp, c_int, _f, _f, c_void_p]),
    'hcgs_lstm_fwd': (c_int, [c_void_p, _f, _f, _f, c_float, c_int, c_int, c_int, _f, _f, _f, c_int, c_void_p]),
    'hcgs_lstm_bwd': (c_int, [c_void_p, _f, _f, _f, _f, _f, c_float, c_int, c_int, c_int, _f, c_void_p]),
    'hcgs_lstm_fwd_ln': (c_int, [c_void_p, _f, _f, _f, c_float, c_int, c_int, c_int, _f, _f, _f, c_int, _f, _f, c_float,
                                 _f, c_void_p]),
    'hcgs_lstm_bwd_ln': (c_int, [c_void_p, _f, _f, _f, _f, _f, c_float, c_int, c_int, c_int, _f, _f, _f, c_float, _f, _f,
                                 c_void_p]),
    'hcgs_logsoftmax_bwd_fused_onehot': (c_int, [c_void_p, c_void_p, _f, _f, c_int64, c_int, _f, c_int64, _f, c_void_p]),
    'hcgs_nll_err_rows': (c_int, [_f, c_int64, c_int64, c_int, c_void_p, _f, _f, c_void_p]),
    'hcgs_gru_fwd': (c_int, [c_void_p, c_int, _f, _f, _f, c_float, c_int, c_int, c_int, _f, _f, _f, c_void_p]),
    'hcgs_gru_bwd': (c_int, [c_void_p, c_int, _f, _f, _f, _f, _f, c_float, c_int, c_int, c_int, _f, c_void_p]),
    'hcgs_lstm_tc_info': (c_int, [c_void_p, c_int, c_int, POINTER(c_int64)]),
    'hcgs_lstm_tc_pack': (c_int, [c_void_p, c_int, _pp, c_int, _f, _f, c_void_p]),
    'hcgs_lstm_tc_fwd': (c_int, [c_void_p, c_int, c_int, _f, _f, c_int, _f, c_float, c_int, c_int, c_int, _f, _f, _f,
                                 _f, _f, _f, _pp, _pp, c_void_p]),
    'hcgs_lstm_tc_fwd_split': (c_int, [c_void_p, c_int, c_int, _f, _f, c_int, _f, c_float, c_int, c_int, c_int, _f, _f,
                                       _f, _f, _f, _f, _f, _pp, _pp, c_void_p]),
    'hcgs_lstm_tc_bwd': (c_int, [c_void_p, c_int, c_int, _f, _f, _f, _f, c_int, _f, c_float, c_int, c_int, c_int, _f,
                                 _f, c_void_p]),
    'hcgs_tc_trace_read': (c_int, [POINTER(c_int64)]),
    'hcgs_tc_timeline_read': (c_int, [POINTER(c_int64)]),
    'hcgs_tc_pack_nt': (c_int, [c_void_p, c_int, _pp, c_int, _f, c_void_p]),
    'hcgs_tc_gemm_nt': (c_int, [c_void_p, c_int, _f, _f, c_int64, c_int64, _pp, _f, c_int64, _f, c_void_p]),
    'hcgs_cast_bf16': (c_int, [_f, c_int64, c_int64, c_int, _f, c_int64, c_void_p]),
    'hcgs_tc_pack_nt_split': (c_int, [c_void_p, c_int, _pp, c_int, _f, c_void_p]),
    'hcgs_tc_gemm_nt_split': (c_int, [c_void_p, c_int, _f, _f, c_int64, _f, c_int64, c_int64, _pp, _f, c_int64, _f,
                                      c_void_p]),
    'hcgs_split_bf16': (c_int, [_f, c_int64, c_int64, c_int, _f, _f, c_int64, c_void_p]),
    'hcgs_tc_pack_nn': (c_int, [c_void_p, c_int, _pp, c_int, _f, c_void_p]),
    'hcgs_tc_gemm_nn': (c_int, [c_void_p, c_int, _f, _f, c_int64, c_int64, _f, c_int64, c_void_p]),
    'hcgs_tc_sddmm': (c_int, [c_void_p, c_int, _f, c_int64, _f, c_int64, c_int64, _pp, c_int64, _f, c_void_p]),
    'hcgs_out_info': (c_int, [c_void_p, POINTER(c_int64)]),
    'hcgs_out_pack': (c_int, [c_void_p, _f, c_int, _f, c_void_p]),
    'hcgs_out_logsoftmax_prior_fwd': (c_int, [c_void_p, _f, _f, c_int64, c_int64, _f, _f, _f, c_int64, c_void_p]),
    'hcgs_act_fwd': (c_int, [_f, _f, _f, c_int64, c_int, c_void_p]),
    'hcgs_act_bwd': (c_int, [_f, _f, _f, _f, c_int64, c_int, c_void_p]),
    'hcgs_logsoftmax_fwd': (c_int, [_f, c_int, _f, c_int, c_int64, c_int, _f, c_void_p]),
    'hcgs_logsoftmax_bwd': (c_int, [_f, _f, _f, c_int64, c_int, c_void_p]),
    'hcgs_logsoftmax_bwd_fused': (c_int, [_f, _f, c_int64, c_int, _f, c_int64, _f, c_void_p]),
    'hcgs_nll_fwd_bwd': (c_int, [_f, c_int64, c_int, _f, c_int, c_float, _f, _f, c_void_p]),
    'hcgs_opt_step': (c_int, [c_int, _f, _f, _f, _f, c_int64, POINTER(c_float), c_float, c_void_p]),
    'hcgs_opt_step_multi': (c_int, [c_int, c_int, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p),
                                    POINTER(c_void_p), POINTER(c_int64), POINTER(ctypes.c_double), c_float, c_void_p]),
    'hcgs_assemble_batch': (c_int, [_f, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, _f, c_void_p]),
    'hcgs_unpad_rows': (c_int, [_f, c_int64, c_int, c_void_p, c_void_p, c_int, c_int, _f, c_void_p]),
    'hcgs_window_abs_mean': (c_int, [_f, c_int64, c_void_p, c_int, _f, c_void_p]),
    'hcgs_fill_rects': (c_int, [_f, c_int64, c_void_p, c_int, c_float, c_void_p]),
}


def load():
    """dlopen the library once; raise NativeError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeError(
            'native library %s is missing: run `python -c "import __graft_entry__ as g; g.build()"` '
            '(or ./build_native.sh) -- there is no CPU fallback' % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.hcgs_abi_version() != 1:
        raise NativeError('ABI version mismatch: library %d, binding 1' % lib.hcgs_abi_version())
    _lib = lib
    return lib


# kernels launched by each entry point (our own kernels only; memsets are not counted)
KERNELS_PER_CALL = {'hcgs_mask_apply': 1, 'hcgs_quantize_w': 1, 'hcgs_quantize_act': 2, 'hcgs_bsgemm_nt': 1,
                    'hcgs_bsgemm_nn': 1, 'hcgs_bs_sddmm': 1, 'hcgs_colsum': 1, 'hcgs_bn_fwd': 3, 'hcgs_bn_fwd_static': 3, 'hcgs_bn_bwd': 2, 'hcgs_bn_fwd_il': 3, 'hcgs_bn_bwd_il': 2, 'hcgs_bn_stats_il': 1,
                    'hcgs_lstm_pack': 2, 'hcgs_lstm_fwd': 1, 'hcgs_lstm_bwd': 1, 'hcgs_lstm_fwd_ln': 1, 'hcgs_lstm_bwd_ln': 1, 'hcgs_gru_fwd': 1, 'hcgs_gru_bwd': 1,
                    'hcgs_lstm_tc_pack': 2, 'hcgs_lstm_tc_fwd': 1, 'hcgs_lstm_tc_fwd_split': 1, 'hcgs_lstm_tc_bwd': 1,
                    'hcgs_tc_pack_nt': 1, 'hcgs_tc_gemm_nt': 1, 'hcgs_cast_bf16': 1,
                    'hcgs_tc_pack_nt_split': 1, 'hcgs_tc_gemm_nt_split': 1, 'hcgs_split_bf16': 1,
                    'hcgs_tc_pack_nn': 1, 'hcgs_tc_gemm_nn': 1, 'hcgs_tc_sddmm': 1, 'hcgs_out_pack': 1,
                    'hcgs_out_logsoftmax_prior_fwd': 1, 'hcgs_act_fwd': 1, 'hcgs_act_bwd': 1,
                    'hcgs_logsoftmax_fwd': 1, 'hcgs_logsoftmax_bwd': 1, 'hcgs_logsoftmax_bwd_fused': 1,
                    'hcgs_layernorm_fwd': 1, 'hcgs_layernorm_bwd': 1, 'hcgs_nll_fwd_bwd': 1, 'hcgs_opt_step': 1,
                    'hcgs_opt_step_multi': 1, 'hcgs_assemble_batch': 1, 'hcgs_unpad_rows': 1, 'hcgs_logsoftmax_bwd_fused_onehot': 1, 'hcgs_nll_err_rows': 1, 'hcgs_window_abs_mean': 1, 'hcgs_fill_rects': 1}
LAUNCHES = {'total': 0}


def reset_launch_count():
    LAUNCHES.clear()
    LAUNCHES['total'] = 0


def _check(rc, what):
    n = KERNELS_PER_CALL.get(what, 0)
    if n:
        LAUNCHES['total'] += n
        LAUNCHES[what] = LAUNCHES.get(what, 0) + n
    if rc != 0:
        msg = load().hcgs_last_error()
        raise NativeError('%s failed (%d): %s' % (what, rc, msg.decode() if msg else '?'))


def _require_cuda(*tensors):
    if not torch.cuda.is_available():
        raise NativeError('no CUDA device: the HCGS hot path has no CPU fallback')
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise NativeError('expected a CUDA tensor, got %s' % t.device)
        if t.dtype != torch.float32:
            raise NativeError('expected float32, got %s' % t.dtype)


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return c_void_p(0) if t is None else c_void_p(t.data_ptr())


def _ptrs(ts):
    """Host array of device pointers (NULL-able entries)."""
    if ts is None:
        return None
    arr = (c_void_p * len(ts))()
    for i, t in enumerate(ts):
        arr[i] = None if t is None else t.data_ptr()
    return ctypes.cast(arr, _pp)


def _row_major_2d(t, what):
    """(tensor, ld): accept any 2-D view whose last dim is contiguous (e.g. a column slice)."""
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1):
        raise NativeError('%s must be a 2-D row-major view, got shape %s stride %s' % (what, tuple(t.shape), t.stride()))
    ld = t.stride(0) if t.shape[0] > 1 else max(t.shape[1], t.stride(0))
    return ld


def device_info(device=0):
    info = (c_int * 8)()
    _check(load().hcgs_device_info(device, info), 'hcgs_device_info')
    return {'sm_count': info[0], 'cc': (info[1], info[2]), 'smem_optin': info[3], 'cluster_launch': bool(info[4])}


class Plan:
    """Owns one ``hcgs_plan`` (device copies of the block index of one mask)."""

    def __init__(self, mask, recurrent=False, cluster_size=0):
        lib = load()
        if not torch.cuda.is_available():
            raise NativeError('no CUDA device: cannot create a plan')
        m = mask.detach().to('cpu', torch.float32).contiguous()
        self.n_out, self.n_in = int(m.shape[0]), int(m.shape[1])
        h = c_void_p()
        _check(lib.hcgs_plan_create(ctypes.cast(m.data_ptr(), POINTER(c_float)), self.n_out, self.n_in,
                                    1 if recurrent else 0, int(cluster_size), ctypes.byref(h)), 'hcgs_plan_create')
        self._h = h
        info = (c_int64 * 16)()
        _check(lib.hcgs_plan_info(self._h, info), 'hcgs_plan_info')
        self.nz_tiles, self.total_tiles, self.nnz = int(info[2]), int(info[3]), int(info[4])
        self.recurrent = bool(recurrent)
        self.cluster_size, self.rows_per_cta, self.bell_width = int(info[5]), int(info[6]), int(info[7])
        self.fwd_pack_floats, self.bwd_pack_floats = int(info[8]), int(info[9])
        self.smem_fwd, self.smem_bwd, self.threads = int(info[10]), int(info[11]), int(info[12])
        self.n_blocks, self.w_in_smem = int(info[13]), bool(info[14])

    @property
    def handle(self):
        if self._h is None:
            raise NativeError('plan already destroyed')
        return self._h

    def destroy(self):
        if getattr(self, '_h', None) is not None and _lib is not None:
            _lib.hcgs_plan_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


# Plans shared across model instances: core.run_nn rebuilds every net (and with it every mask-derived plan) for each
# chunk (core.py:107-121).  The masks of a run come from one seeded generator or from the checkpoint, so chunk after
# chunk they are the same bits: the plan -- block index, cluster layout, task tables -- is looked up by a digest of the
# structural mask instead of being rebuilt.  Entries are only referenced, never destroyed, here: a plan
# is freed by Plan.__del__ once the store and every net have dropped it.
_PLAN_STORE = collections.OrderedDict()
PLAN_STORE_LIMIT = 256
PLAN_STORE_STATS = {'hits': 0, 'misses': 0}


def shared_plan(mask, recurrent=False):
    m = mask.detach().to('cpu', torch.float32).contiguous()
    nz = (m != 0).numpy()
    key = (tuple(m.shape), bool(recurrent), torch.cuda.current_device() if torch.cuda.is_available() else -1,
           hashlib.blake2b(np.packbits(nz).tobytes(), digest_size=16).digest())
    plan = _PLAN_STORE.get(key)
    if plan is not None and plan._h is not None:
        _PLAN_STORE.move_to_end(key)
        PLAN_STORE_STATS['hits'] += 1
        return plan
    PLAN_STORE_STATS['misses'] += 1
    plan = Plan(m, recurrent=recurrent)
    _PLAN_STORE[key] = plan
    while len(_PLAN_STORE) > PLAN_STORE_LIMIT:
        _PLAN_STORE.popitem(last=False)
    return plan


# ---------------------------------------------------------------------------------------------
# thin typed wrappers (argument checking only; all arithmetic is in the library)
# ---------------------------------------------------------------------------------------------

def mask_apply(weights, mask):
    _require_cuda(mask, *weights)
    n = mask.numel()
    for w in weights:
        assert w.is_contiguous() and w.numel() == n
    assert mask.is_contiguous()
    _check(load().hcgs_mask_apply(_ptrs(weights), len(weights), _ptr(mask), n, _stream()), 'hcgs_mask_apply')


def quantize_w(master, out, bits):
    _require_cuda(master, out)
    assert master.is_contiguous() and out.is_contiguous() and master.numel() == out.numel()
    _check(load().hcgs_quantize_w(_ptr(master), _ptr(out), master.numel(), int(bits), _stream()), 'hcgs_quantize_w')


def quantize_act(x2d, y2d, bits, repeat, scratch):
    _require_cuda(x2d, y2d, scratch)
    ldx, ldy = _row_major_2d(x2d, 'x'), _row_major_2d(y2d, 'y')
    _check(load().hcgs_quantize_act(_ptr(x2d), ldx, _ptr(y2d), ldy, x2d.shape[0], x2d.shape[1], int(bits),
                                    int(repeat), _ptr(scratch), _stream()), 'hcgs_quantize_act')


def bsgemm_nt(plan, weights, x2d, y2d, biases=None):
    """y2d[:, g*n_out:(g+1)*n_out] = x2d @ weights[g].T (+ bias[g])"""
    _require_cuda(x2d, y2d, *weights)
    ldw = _row_major_2d(weights[0], 'w')
    for w in weights:
        assert tuple(w.shape) == (plan.n_out, plan.n_in) and _row_major_2d(w, 'w') == ldw
    assert x2d.shape[1] == plan.n_in and y2d.shape[1] == len(weights) * plan.n_out and y2d.shape[0] == x2d.shape[0]
    _check(load().hcgs_bsgemm_nt(plan.handle, len(weights), _ptrs(weights), ldw, _ptr(x2d),
                                 _row_major_2d(x2d, 'x'), x2d.shape[0], _ptrs(biases), _ptr(y2d),
                                 _row_major_2d(y2d, 'y'), _stream()), 'hcgs_bsgemm_nt')


def bsgemm_nn(plan, weights, dy2d, dx2d):
    _require_cuda(dy2d, dx2d, *weights)
    ldw = _row_major_2d(weights[0], 'w')
    assert dy2d.shape[1] == len(weights) * plan.n_out and dx2d.shape[1] == plan.n_in
    _check(load().hcgs_bsgemm_nn(plan.handle, len(weights), _ptrs(weights), ldw, _ptr(dy2d),
                                 _row_major_2d(dy2d, 'dy'), dy2d.shape[0], _ptr(dx2d),
                                 _row_major_2d(dx2d, 'dx'), _stream()), 'hcgs_bsgemm_nn')


def bs_sddmm(plan, dy2d, x2d, dws, mask=None):
    _require_cuda(dy2d, x2d, mask, *dws)
    lddw = _row_major_2d(dws[0], 'dw')
    assert dy2d.shape[0] == x2d.shape[0] and x2d.shape[1] == plan.n_in
    _check(load().hcgs_bs_sddmm(plan.handle, len(dws), _ptr(dy2d), _row_major_2d(dy2d, 'dy'), _ptr(x2d),
                                _row_major_2d(x2d, 'x'), x2d.shape[0], _ptrs(dws), lddw, _ptr(mask), _stream()),
           'hcgs_bs_sddmm')


def colsum(dy2d, n_groups, group_cols, outs):
    _require_cuda(dy2d, *outs)
    _check(load().hcgs_colsum(_ptr(dy2d), _row_major_2d(dy2d, 'dy'), dy2d.shape[0], n_groups, group_cols,
                              _ptrs(outs), _stream()), 'hcgs_colsum')


def bn_fwd(x2d, y2d, n_groups, group_cols, gammas, betas, rmeans, rvars, eps, momentum, training, save_mean,
           save_invstd, workspace):
    _require_cuda(x2d, y2d, save_mean, save_invstd, workspace)
    _check(load().hcgs_bn_fwd(_ptr(x2d), _row_major_2d(x2d, 'x'), _ptr(y2d), _row_major_2d(y2d, 'y'), x2d.shape[0],
                              n_groups, group_cols, _ptrs(gammas), _ptrs(betas), _ptrs(rmeans), _ptrs(rvars),
                              float(eps), float(momentum), 1 if training else 0, _ptr(save_mean),
                              _ptr(save_invstd), _ptr(workspace),
                              0 if workspace is None else workspace.numel(), _stream()), 'hcgs_bn_fwd')


BN_STATIC_MAX_COLS = 4096   # HCGS_BN_STATIC_MAX_COLS of include/hcgs_b200.h


def bn_fwd_static(x2d, y2d, n_groups, group_cols, gammas, betas, rmeans, rvars, eps, momentum, training, save_mean,
                  save_invstd):
    """hcgs_bn_fwd without a workspace: the scratch is a static device array of the library (no allocation here or
    in the library; n_groups * group_cols <= BN_STATIC_MAX_COLS; overlapping calls race on the shared scratch)."""
    _require_cuda(x2d, y2d, save_mean, save_invstd)
    _check(load().hcgs_bn_fwd_static(_ptr(x2d), _row_major_2d(x2d, 'x'), _ptr(y2d), _row_major_2d(y2d, 'y'),
                                     x2d.shape[0], n_groups, group_cols, _ptrs(gammas), _ptrs(betas), _ptrs(rmeans),
                                     _ptrs(rvars), float(eps), float(momentum), 1 if training else 0,
                                     _ptr(save_mean), _ptr(save_invstd), _stream()), 'hcgs_bn_fwd_static')


def bn_bwd(dy2d, x2d, n_groups, group_cols, gammas, save_mean, save_invstd, dx2d, dgammas, dbetas, workspace):
    _require_cuda(dy2d, x2d, dx2d, save_mean, save_invstd, workspace)
    _check(load().hcgs_bn_bwd(_ptr(dy2d), _row_major_2d(dy2d, 'dy'), _ptr(x2d), _row_major_2d(x2d, 'x'),
                              x2d.shape[0], n_groups, group_cols, _ptrs(gammas), _ptr(save_mean),
                              _ptr(save_invstd), _ptr(dx2d), _row_major_2d(dx2d, 'dx'), _ptrs(dgammas),
                              _ptrs(dbetas), _ptr(workspace), workspace.numel(), _stream()), 'hcgs_bn_bwd')


def bn_fwd_il(x2d, y2d, n_groups, units, ld_units, gammas, betas, rmeans, rvars, eps, momentum, training,
              save_mean, save_invstd, workspace):
    """gate-interleaved BN over [m, n_groups*ld_units] (tensor-core build); y2d may alias x2d"""
    _require_cuda(x2d, y2d, save_mean, save_invstd, workspace)
    assert x2d.is_contiguous() and y2d.is_contiguous() and x2d.shape[1] == n_groups * ld_units
    _check(load().hcgs_bn_fwd_il(_ptr(x2d), _ptr(y2d), x2d.shape[0], n_groups, units, ld_units, _ptrs(gammas),
                                 _ptrs(betas), _ptrs(rmeans), _ptrs(rvars), float(eps), float(momentum),
                                 1 if training else 0, _ptr(save_mean), _ptr(save_invstd), _ptr(workspace),
                                 0 if workspace is None else workspace.numel(), _stream()), 'hcgs_bn_fwd_il')


def bn_bwd_il(dy2d, x2d, n_groups, units, ld_units, gammas, save_mean, save_invstd, dx2d, dx_bf16, dgammas, dbetas,
              workspace):
    """dy2d: fp32 or bf16 [m, n_groups*ld_units] (the bf16 form is what lstm_tc_bwd writes for the GEMMs)"""
    _require_cuda(x2d, dx2d, save_mean, save_invstd, workspace)
    _require_cuda_any(dx_bf16, dy2d)
    assert dy2d.is_contiguous() and x2d.is_contiguous() and dy2d.shape[1] == n_groups * ld_units
    dy_f, dy_b = (dy2d, None) if dy2d.dtype == torch.float32 else (None, dy2d)
    assert dy2d.dtype in (torch.float32, torch.bfloat16)
    _check(load().hcgs_bn_bwd_il(_ptr(dy_f), _ptr(dy_b), _ptr(x2d), x2d.shape[0], n_groups, units, ld_units, _ptrs(gammas),
                                 _ptr(save_mean), _ptr(save_invstd), _ptr(dx2d), _ptr(dx_bf16), _ptrs(dgammas),
                                 _ptrs(dbetas), _ptr(workspace), workspace.numel(), _stream()), 'hcgs_bn_bwd_il')


def bn_stats_il(partial, rows_per_split, m, n_groups, units, ld_units, rmeans, rvars, eps, momentum, training,
                save_mean, save_invstd):
    """statistics only (no apply pass): merge the GEMM epilogue's per-tile partials (training) or hand over the
    running statistics (eval) as save_mean / save_invstd for the fused normalisation in lstm_tc_fwd"""
    _require_cuda(partial, save_mean, save_invstd)
    splits = 0 if partial is None else partial.shape[0]
    _check(load().hcgs_bn_stats_il(_ptr(partial), splits, int(rows_per_split), int(m), n_groups, units, ld_units,
                                   _ptrs(rmeans), _ptrs(rvars), float(eps), float(momentum), 1 if training else 0,
                                   _ptr(save_mean), _ptr(save_invstd), _stream()), 'hcgs_bn_stats_il')


def layernorm_fwd(x2d, y2d, gamma, beta, eps, save_mean=None, save_std=None):
    """reference LayerNorm over the last dimension (unbiased std, divide by std + eps)"""
    _require_cuda(x2d, y2d, gamma, beta, save_mean, save_std)
    _check(load().hcgs_layernorm_fwd(_ptr(x2d), _row_major_2d(x2d, 'x'), _ptr(y2d), _row_major_2d(y2d, 'y'),
                                     x2d.shape[0], x2d.shape[1], _ptr(gamma), _ptr(beta), float(eps), _ptr(save_mean),
                                     _ptr(save_std), _stream()), 'hcgs_layernorm_fwd')


def layernorm_bwd(dy2d, x2d, gamma, eps, save_mean, save_std, dx2d, dgamma, dbeta):
    """dx written (or skipped when None); dgamma / dbeta are ADDED to (zero-initialise them)"""
    _require_cuda(dy2d, x2d, gamma, save_mean, save_std, dx2d, dgamma, dbeta)
    _check(load().hcgs_layernorm_bwd(_ptr(dy2d), _row_major_2d(dy2d, 'dy'), _ptr(x2d), _row_major_2d(x2d, 'x'),
                                     x2d.shape[0], x2d.shape[1], _ptr(gamma), float(eps), _ptr(save_mean),
                                     _ptr(save_std), _ptr(dx2d), _row_major_2d(dx2d, 'dx') if dx2d is not None else 0,
                                     _ptr(dgamma), _ptr(dbeta), _stream()), 'hcgs_layernorm_bwd')


def lstm_pack(plan, us, fwd_pack=None, bwd_pack=None):
    _require_cuda(fwd_pack, bwd_pack, *us)
    ldu = _row_major_2d(us[0], 'u')
    assert len(us) == 4
    _check(load().hcgs_lstm_pack(plan.handle, _ptrs(us), ldu, _ptr(fwd_pack), _ptr(bwd_pack), _stream()),
           'hcgs_lstm_pack')


def lstm_fwd(plan, fwd_pack, pre, drop, drop_scalar, act, h_out, c_out=None, gates_out=None, act_bits=0):
    _require_cuda(fwd_pack, pre, drop, h_out, c_out, gates_out)
    T, B = pre.shape[0], pre.shape[1]
    assert pre.is_contiguous() and h_out.is_contiguous() and pre.shape[2] == 4 * plan.n_out
    _check(load().hcgs_lstm_fwd(plan.handle, _ptr(fwd_pack), _ptr(pre), _ptr(drop), float(drop_scalar), int(act),
                                T, B, _ptr(h_out), _ptr(c_out), _ptr(gates_out), int(act_bits), _stream()),
           'hcgs_lstm_fwd')


def lstm_bwd(plan, bwd_pack, dh_out, gates, c, drop, drop_scalar, act, dpre):
    _require_cuda(bwd_pack, dh_out, gates, c, drop, dpre)
    T, B = dh_out.shape[0], dh_out.shape[1]
    assert dh_out.is_contiguous() and gates.is_contiguous() and c.is_contiguous() and dpre.is_contiguous()
    _check(load().hcgs_lstm_bwd(plan.handle, _ptr(bwd_pack), _ptr(dh_out), _ptr(gates), _ptr(c), _ptr(drop),
                                float(drop_scalar), int(act), T, B, _ptr(dpre), _stream()), 'hcgs_lstm_bwd')


def lstm_fwd_ln(plan, fwd_pack, pre, drop, drop_scalar, act, h_out, c_out, gates_out, act_bits, gamma, beta, eps,
                stats=None):
    """lstm_fwd with the reference's LayerNorm of h_t inside the time loop; stats [T,B,2] = {mean, std + eps}."""
    _require_cuda(fwd_pack, pre, drop, h_out, c_out, gates_out, gamma, beta, stats)
    T, B = pre.shape[0], pre.shape[1]
    H = plan.n_out
    assert pre.is_contiguous() and h_out.is_contiguous() and pre.shape[2] == 4 * H
    assert gamma.is_contiguous() and beta.is_contiguous() and gamma.numel() == H and beta.numel() == H
    for t, shape in ((c_out, (T, B, H)), (gates_out, (T, B, 4 * H)), (stats, (T, B, 2))):
        assert t is None or (t.is_contiguous() and tuple(t.shape) == shape)
    _check(load().hcgs_lstm_fwd_ln(plan.handle, _ptr(fwd_pack), _ptr(pre), _ptr(drop), float(drop_scalar), int(act), T, B,
                                   _ptr(h_out), _ptr(c_out), _ptr(gates_out), int(act_bits), _ptr(gamma), _ptr(beta),
                                   float(eps), _ptr(stats), _stream()), 'hcgs_lstm_fwd_ln')


def lstm_bwd_ln(plan, bwd_pack, dh_out, gates, c, drop, drop_scalar, act, dpre, gamma, stats, eps):
    """lstm_bwd through the in-loop LayerNorm; returns (dgamma, dbeta) [H]."""
    _require_cuda(bwd_pack, dh_out, gates, c, drop, dpre, gamma, stats)
    T, B = dh_out.shape[0], dh_out.shape[1]
    H = plan.n_out
    assert dh_out.is_contiguous() and gates.is_contiguous() and c.is_contiguous() and dpre.is_contiguous()
    assert stats.is_contiguous() and tuple(stats.shape) == (T, B, 2) and gamma.is_contiguous() and gamma.numel() == H
    groups = (B + 7) // 8
    dg = torch.empty(groups, H, device=dpre.device, dtype=torch.float32)
    db = torch.empty(groups, H, device=dpre.device, dtype=torch.float32)
    _check(load().hcgs_lstm_bwd_ln(plan.handle, _ptr(bwd_pack), _ptr(dh_out), _ptr(gates), _ptr(c), _ptr(drop),
                                   float(drop_scalar), int(act), T, B, _ptr(dpre), _ptr(gamma), _ptr(stats), float(eps),
                                   _ptr(dg), _ptr(db), _stream()), 'hcgs_lstm_bwd_ln')
    return dg.sum(0), db.sum(0)


CELL_CODES = {'gru': 0, 'minimalgru': 1, 'ligru': 2, 'rnn': 3}
CELL_GATES = {'gru': ('z', 'r', 'h'), 'minimalgru': ('z', 'h'), 'ligru': ('z', 'h'), 'rnn': ('h',)}


def gru_pack(plan, cell, us, fwd_pack=None, bwd_pack=None):
    """us: dict gate letter -> U [H,H] for the gates of `cell`; packed into the LSTM's four-slot layout
    (slot 0 = update, 1 = reset, 2 = candidate; absent gates get a placeholder that is never read)."""
    uh = us['h']
    lstm_pack(plan, [us.get('z', uh), us.get('r', uh), uh, uh], fwd_pack, bwd_pack)


def gru_fwd(plan, cell, fwd_pack, pre, drop, drop_scalar, act, h_out, m_out=None, gates_out=None):
    """pre / gates_out [T,B,NG*H] (gate-major: z, r, candidate as far as the cell has them), h_out / m_out [T,B,H]"""
    _require_cuda(fwd_pack, pre, drop, h_out, m_out, gates_out)
    T, B = pre.shape[0], pre.shape[1]
    ng = len(CELL_GATES[cell])
    assert pre.is_contiguous() and h_out.is_contiguous() and pre.shape[2] == ng * plan.n_out
    for t in (m_out, gates_out):
        assert t is None or t.is_contiguous()
    _check(load().hcgs_gru_fwd(plan.handle, CELL_CODES[cell], _ptr(fwd_pack), _ptr(pre), _ptr(drop), float(drop_scalar),
                               int(act), T, B, _ptr(h_out), _ptr(m_out), _ptr(gates_out), _stream()), 'hcgs_gru_fwd')


def gru_bwd(plan, cell, bwd_pack, dh_out, gates, h, drop, drop_scalar, act, dpre):
    _require_cuda(bwd_pack, dh_out, gates, h, drop, dpre)
    T, B = dh_out.shape[0], dh_out.shape[1]
    assert dh_out.is_contiguous() and gates.is_contiguous() and h.is_contiguous() and dpre.is_contiguous()
    assert dpre.shape == gates.shape and gates.shape[2] == len(CELL_GATES[cell]) * plan.n_out
    _check(load().hcgs_gru_bwd(plan.handle, CELL_CODES[cell], _ptr(bwd_pack), _ptr(dh_out), _ptr(gates), _ptr(h),
                               _ptr(drop), float(drop_scalar), int(act), T, B, _ptr(dpre), _stream()), 'hcgs_gru_bwd')


def _require_cuda_any(*tensors):
    if not torch.cuda.is_available():
        raise NativeError('no CUDA device: the HCGS hot path has no CPU fallback')
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NativeError('expected a CUDA tensor, got %s' % t.device)


class TcConfig:
    """Resolved launch shape of the tensor-core recurrence for one plan (hcgs_lstm_tc_info)."""

    def __init__(self, plan, cluster_size=0, nb=0):
        info = (c_int64 * 16)()
        _check(load().hcgs_lstm_tc_info(plan.handle, int(cluster_size), int(nb), info), 'hcgs_lstm_tc_info')
        self.cluster_size, self.nb = int(info[0]), int(info[1])
        self.fwd_pack_u32, self.bwd_pack_u32 = int(info[2]), int(info[3])
        self.smem_fwd, self.smem_bwd = int(info[4]), int(info[5])
        self.warps, self.strips_per_cta, self.k_steps = int(info[6]), int(info[7]), int(info[8])
        self.n_blocks, self.strips_per_warp, self.padded_hidden = int(info[9]), int(info[10]), int(info[11])
        self.max_clusters_fwd, self.max_clusters_bwd = int(info[12]), int(info[13])


def lstm_tc_pack(plan, cfg, us, fwd_pack=None, bwd_pack=None):
    """fwd_pack / bwd_pack: int32 CUDA tensors of cfg.fwd_pack_u32 / cfg.bwd_pack_u32 elements."""
    _require_cuda(*us)
    _require_cuda_any(fwd_pack, bwd_pack)
    assert len(us) == 4
    _check(load().hcgs_lstm_tc_pack(plan.handle, cfg.cluster_size, _ptrs(us), _row_major_2d(us[0], 'u'),
                                    _ptr(fwd_pack), _ptr(bwd_pack), _stream()), 'hcgs_lstm_tc_pack')


def pad8(n):
    return (n + 7) // 8 * 8


def lstm_tc_fwd(plan, cfg, fwd_pack, pre, drop, drop_scalar, act, T, B, h_out=None, hb_out=None, c_out=None,
                gates_out=None, bn=None, hl_out=None):
    """pre / gates_out: [T*B, 4*ldH] fp32 gate-interleaved (column u*4+g), ldH = pad8(H);
    c_out / h_out [T*B, ldH] fp32, hb_out [T*B, ldH] bf16, pad columns zero (all optional but one of h/hb).
    bn = (mean, invstd, gammas, betas): batch norm of `pre` fused into the kernel (bn_stats_il)."""
    _require_cuda(pre, drop, h_out, c_out, gates_out)
    bn_mean = bn_invstd = bn_g = bn_b = None
    if bn is not None:
        bn_mean, bn_invstd, bn_g, bn_b = bn
        _require_cuda(bn_mean, bn_invstd, *bn_g, *bn_b)
        assert bn_mean.numel() == pre.shape[-1] and bn_invstd.numel() == pre.shape[-1] and len(bn_g) == 4
    _require_cuda_any(fwd_pack, hb_out)
    H = plan.n_out
    ldH = pre.shape[-1] // 4
    assert pre.is_contiguous() and pre.numel() == T * B * 4 * ldH and ldH % 8 == 0 and ldH >= H
    for t, n in ((h_out, ldH), (hb_out, ldH), (c_out, ldH), (gates_out, 4 * ldH)):
        assert t is None or (t.is_contiguous() and t.numel() == T * B * n)
    if hl_out is not None:
        # hl_out [T*B, ldH] bf16 = bf16(h - float(bf16(h))): the low term of the next layer's two-term projection
        _require_cuda_any(hl_out)
        assert hb_out is not None and hl_out.is_contiguous() and hl_out.numel() == T * B * ldH
        _check(load().hcgs_lstm_tc_fwd_split(plan.handle, cfg.cluster_size, cfg.nb, _ptr(fwd_pack), _ptr(pre), ldH,
                                             _ptr(drop), float(drop_scalar), int(act), T, B, _ptr(h_out), _ptr(hb_out),
                                             _ptr(hl_out), _ptr(c_out), _ptr(gates_out), _ptr(bn_mean), _ptr(bn_invstd),
                                             _ptrs(bn_g), _ptrs(bn_b), _stream()), 'hcgs_lstm_tc_fwd_split')
        return
    _check(load().hcgs_lstm_tc_fwd(plan.handle, cfg.cluster_size, cfg.nb, _ptr(fwd_pack), _ptr(pre), ldH, _ptr(drop),
                                   float(drop_scalar), int(act), T, B, _ptr(h_out), _ptr(hb_out), _ptr(c_out),
                                   _ptr(gates_out), _ptr(bn_mean), _ptr(bn_invstd), _ptrs(bn_g), _ptrs(bn_b),
                                   _stream()), 'hcgs_lstm_tc_fwd')


def lstm_tc_bwd(plan, cfg, bwd_pack, dh_out, gates, c, drop, drop_scalar, act, T, B, dpre, dpre_bf16=None):
    """gates / dpre / dpre_bf16: [T*B, 4*ldH] gate-interleaved; c and dh_out [T*B, ldH] (pad columns ignored).
    Either of dpre (fp32) / dpre_bf16 may be None, not both."""
    _require_cuda(dh_out, gates, c, drop, dpre)
    _require_cuda_any(bwd_pack, dpre_bf16)
    assert dpre is not None or dpre_bf16 is not None
    H = plan.n_out
    ldH = gates.shape[-1] // 4
    assert dh_out.is_contiguous() and gates.is_contiguous() and c.is_contiguous()
    assert dh_out.numel() == T * B * ldH and gates.numel() == T * B * 4 * ldH and c.numel() == T * B * ldH
    for d in (dpre, dpre_bf16):
        assert d is None or (d.is_contiguous() and d.numel() == gates.numel())
    _check(load().hcgs_lstm_tc_bwd(plan.handle, cfg.cluster_size, cfg.nb, _ptr(bwd_pack), _ptr(dh_out), _ptr(gates),
                                   _ptr(c), ldH, _ptr(drop), float(drop_scalar), int(act), T, B, _ptr(dpre),
                                   _ptr(dpre_bf16), _stream()), 'hcgs_lstm_tc_bwd')


def tc_trace_read():
    """per-phase clock totals of the last recurrent launch made with HCGS_TC_DEBUG=64 (tuning aid)"""
    out = (c_int64 * 16)()
    _check(load().hcgs_tc_trace_read(out), 'hcgs_tc_trace_read')
    return [int(v) for v in out]


def tc_timeline_read():
    """clock stamps [warp][16] of CTA 0 at the middle time step of the last hcgs_lstm_tc_bwd made with HCGS_TC_DEBUG=128"""
    out = (c_int64 * 320)()
    _check(load().hcgs_tc_timeline_read(out), 'hcgs_tc_timeline_read')
    return [[int(out[w * 16 + i]) for i in range(16)] for w in range(20)]


def tc_pack_nt(plan, weights, wpack):
    """wpack: bf16 CUDA tensor of plan.nz_tiles * 32*len(weights) * 32 elements."""
    _require_cuda(*weights)
    _require_cuda_any(wpack)
    assert wpack.dtype == torch.bfloat16 and wpack.numel() >= plan.nz_tiles * 32 * len(weights) * 32
    _check(load().hcgs_tc_pack_nt(plan.handle, len(weights), _ptrs(weights), _row_major_2d(weights[0], 'w'),
                                  _ptr(wpack), _stream()), 'hcgs_tc_pack_nt')


def tc_gemm_nt(plan, n_w, wpack, x_bf16, y, biases=None, stats=None):
    """y[:, (u, g)] = x @ W_g^T ; x_bf16 [m, ldx] bf16 (pad columns zero), y [m, ldy] fp32.
    stats (n_w = 4): fp32 [ceil(m/128), ldy, 2] per-tile column {mean, M2} for bn_stats_il."""
    _require_cuda(y, stats)
    assert stats is None or (stats.is_contiguous() and stats.numel() == (x_bf16.shape[0] + 127) // 128 * y.stride(0) * 2)
    _require_cuda_any(wpack, x_bf16)
    assert x_bf16.dtype == torch.bfloat16 and x_bf16.dim() == 2 and x_bf16.stride(1) == 1
    _check(load().hcgs_tc_gemm_nt(plan.handle, int(n_w), _ptr(wpack), _ptr(x_bf16), x_bf16.stride(0), x_bf16.shape[0],
                                  _ptrs(biases), _ptr(y), y.stride(0), _ptr(stats), _stream()), 'hcgs_tc_gemm_nt')


def tc_pack_nt_split(plan, weights, wpack2):
    """wpack2: bf16 CUDA tensor of 2 * plan.nz_tiles * 32*len(weights) * 32 elements ([hi | lo] per kept tile)."""
    _require_cuda(*weights)
    _require_cuda_any(wpack2)
    assert wpack2.dtype == torch.bfloat16 and wpack2.numel() >= 2 * plan.nz_tiles * 32 * len(weights) * 32
    _check(load().hcgs_tc_pack_nt_split(plan.handle, len(weights), _ptrs(weights), _row_major_2d(weights[0], 'w'),
                                        _ptr(wpack2), _stream()), 'hcgs_tc_pack_nt_split')


def tc_gemm_nt_split(plan, n_w, wpack2, x_hi, x_lo, y, biases=None, stats=None):
    """tc_gemm_nt with two-term operands: y = x_hi w_hi^T + x_lo w_hi^T + x_hi w_lo^T (fp32 accumulate)."""
    _require_cuda(y, stats)
    assert stats is None or (stats.is_contiguous() and stats.numel() == (x_hi.shape[0] + 127) // 128 * y.stride(0) * 2)
    _require_cuda_any(wpack2, x_hi, x_lo)
    for t in (x_hi, x_lo):
        assert t.dtype == torch.bfloat16 and t.dim() == 2 and t.stride(1) == 1
    assert x_hi.shape[0] == x_lo.shape[0]
    _check(load().hcgs_tc_gemm_nt_split(plan.handle, int(n_w), _ptr(wpack2), _ptr(x_hi), x_hi.stride(0), _ptr(x_lo),
                                        x_lo.stride(0), x_hi.shape[0], _ptrs(biases), _ptr(y), y.stride(0),
                                        _ptr(stats), _stream()), 'hcgs_tc_gemm_nt_split')


def split_bf16(x2d, hi, lo):
    """fp32 rows -> hi = bf16(x) (hi may be None: only the residual is wanted) and lo = bf16(x - hi)."""
    _require_cuda(x2d)
    _require_cuda_any(hi, lo)
    assert lo.dtype == torch.bfloat16 and lo.shape[0] == x2d.shape[0] and lo.stride(1) == 1
    assert hi is None or (hi.dtype == torch.bfloat16 and hi.shape == lo.shape and hi.stride(0) == lo.stride(0))
    _check(load().hcgs_split_bf16(_ptr(x2d), _row_major_2d(x2d, 'x'), x2d.shape[0], x2d.shape[1], _ptr(hi), _ptr(lo),
                                  lo.stride(0), _stream()), 'hcgs_split_bf16')


def tc_pack_nn(plan, weights, wpack):
    _require_cuda(*weights)
    _require_cuda_any(wpack)
    assert wpack.dtype == torch.bfloat16 and wpack.numel() >= plan.nz_tiles * 32 * len(weights) * 32
    _check(load().hcgs_tc_pack_nn(plan.handle, len(weights), _ptrs(weights), _row_major_2d(weights[0], 'w'),
                                  _ptr(wpack), _stream()), 'hcgs_tc_pack_nn')


def tc_gemm_nn(plan, n_w, wpack, dy_bf16, dx):
    """dx[m, k] = sum_(u,g) dy[m, (u,g)] W_g[u, k]; dy_bf16 [m, lddy] bf16, dx [m, >= n_in] fp32."""
    _require_cuda(dx)
    _require_cuda_any(wpack, dy_bf16)
    assert dy_bf16.dtype == torch.bfloat16 and dy_bf16.dim() == 2 and dy_bf16.stride(1) == 1
    assert dx.shape[0] == dy_bf16.shape[0]
    _check(load().hcgs_tc_gemm_nn(plan.handle, int(n_w), _ptr(wpack), _ptr(dy_bf16), dy_bf16.stride(0),
                                  dy_bf16.shape[0], _ptr(dx), dx.stride(0), _stream()), 'hcgs_tc_gemm_nn')


def tc_sddmm(plan, n_w, dy_bf16, x_bf16, dws, mask=None):
    """dws[g][u, k] += mask * sum_m dy[m, (u,g)] x[m, k]   (dws zero-initialised by the caller)"""
    _require_cuda(mask, *dws)
    _require_cuda_any(dy_bf16, x_bf16)
    assert dy_bf16.dtype == torch.bfloat16 and x_bf16.dtype == torch.bfloat16
    assert dy_bf16.shape[0] == x_bf16.shape[0] and len(dws) == n_w
    _check(load().hcgs_tc_sddmm(plan.handle, int(n_w), _ptr(dy_bf16), dy_bf16.stride(0), _ptr(x_bf16),
                                x_bf16.stride(0), x_bf16.shape[0], _ptrs(dws), _row_major_2d(dws[0], 'dw'),
                                _ptr(mask), _stream()), 'hcgs_tc_sddmm')


def cast_bf16(x2d, out):
    _require_cuda(x2d)
    _require_cuda_any(out)
    assert out.dtype == torch.bfloat16 and out.shape[0] == x2d.shape[0] and out.stride(1) == 1
    _check(load().hcgs_cast_bf16(_ptr(x2d), _row_major_2d(x2d, 'x'), x2d.shape[0], x2d.shape[1], _ptr(out),
                                 out.stride(0), _stream()), 'hcgs_cast_bf16')


def out_info(plan):
    """Shape of the fused output-layer kernel for this plan (cluster 0 = too wide, use gemm + logsoftmax)."""
    if '_out_info' in plan.__dict__:
        return plan.__dict__['_out_info']
    info = (c_int64 * 8)()
    _check(load().hcgs_out_info(plan.handle, info), 'hcgs_out_info')
    plan.__dict__['_out_info'] = {'cluster': int(info[0]), 'strips': int(info[1]), 'tiles': int(info[2]), 'tmem_cols': int(info[3]),
            'smem': int(info[4]), 'items': int(info[5]), 'pack_bytes': int(info[6]),
                                     'stages': int(info[7])}
    return plan.__dict__['_out_info']


def out_pack(plan, w, wpack):
    """wpack: bf16 CUDA tensor of out_info(plan)['pack_bytes'] / 2 elements."""
    _require_cuda(w)
    _require_cuda_any(wpack)
    assert wpack.dtype == torch.bfloat16 and wpack.numel() * 2 >= out_info(plan)['pack_bytes']
    _check(load().hcgs_out_pack(plan.handle, _ptr(w), _row_major_2d(w, 'w'), _ptr(wpack), _stream()), 'hcgs_out_pack')


def out_logsoftmax_prior_fwd(plan, wpack, x_bf16, bias, log_prior, out):
    """out = log_softmax(x W^T + bias) - log_prior, one kernel (logits stay in TMEM)."""
    _require_cuda(bias, log_prior, out)
    _require_cuda_any(wpack, x_bf16)
    assert x_bf16.dtype == torch.bfloat16 and x_bf16.dim() == 2 and x_bf16.stride(1) == 1
    assert out.shape[0] == x_bf16.shape[0] and out.shape[1] == plan.n_out
    _check(load().hcgs_out_logsoftmax_prior_fwd(plan.handle, _ptr(wpack), _ptr(x_bf16), x_bf16.stride(0),
                                                x_bf16.shape[0], _ptr(bias), _ptr(log_prior), _ptr(out),
                                                _row_major_2d(out, 'out'), _stream()),
           'hcgs_out_logsoftmax_prior_fwd')


def act_fwd(x, y, drop_mask, act):
    _require_cuda(x, y, drop_mask)
    assert x.is_contiguous() and y.is_contiguous()
    _check(load().hcgs_act_fwd(_ptr(x), _ptr(y), _ptr(drop_mask), x.numel(), int(act), _stream()), 'hcgs_act_fwd')


def act_bwd(dy, y, drop_mask, dx, act):
    _require_cuda(dy, y, drop_mask, dx)
    assert dy.is_contiguous() and y.is_contiguous() and dx.is_contiguous()
    _check(load().hcgs_act_bwd(_ptr(dy), _ptr(y), _ptr(drop_mask), _ptr(dx), dy.numel(), int(act), _stream()),
           'hcgs_act_bwd')


def logsoftmax_fwd(z2d, out2d, log_prior=None):
    _require_cuda(z2d, out2d, log_prior)
    _check(load().hcgs_logsoftmax_fwd(_ptr(z2d), _row_major_2d(z2d, 'z'), _ptr(out2d), _row_major_2d(out2d, 'out'),
                                      z2d.shape[0], z2d.shape[1], _ptr(log_prior), _stream()), 'hcgs_logsoftmax_fwd')


def logsoftmax_bwd(dlogp, logp, dz):
    _require_cuda(dlogp, logp, dz)
    assert dlogp.is_contiguous() and logp.is_contiguous() and dz.is_contiguous()
    _check(load().hcgs_logsoftmax_bwd(_ptr(dlogp), _ptr(logp), _ptr(dz), logp.shape[0], logp.shape[1], _stream()),
           'hcgs_logsoftmax_bwd')


def logsoftmax_bwd_fused(dlogp, logp, dz_bf16, db=None):
    """dz = dlogp - exp(logp) * rowsum(dlogp) written as bf16 [m, ld] (pad columns zero) and, when db is given,
    db += column sums of the fp32 dz (db zero-initialised by the caller).  Needs p % 4 == 0, p <= 4096."""
    _require_cuda(dlogp, logp, db)
    _require_cuda_any(dz_bf16)
    assert dlogp.is_contiguous() and logp.is_contiguous() and dz_bf16.dtype == torch.bfloat16
    assert dz_bf16.shape[0] == logp.shape[0] and dz_bf16.stride(1) == 1
    _check(load().hcgs_logsoftmax_bwd_fused(_ptr(dlogp), _ptr(logp), logp.shape[0], logp.shape[1], _ptr(dz_bf16),
                                            dz_bf16.stride(0), _ptr(db), _stream()), 'hcgs_logsoftmax_bwd_fused')


def logsoftmax_bwd_fused_onehot(rows, cols, vals, logp, dz_bf16, db=None):
    """logsoftmax_bwd_fused for a gradient with one entry per row (COO rows == arange(m), cols, vals): the dense
    [m, p] gradient of an NLL loss is never materialised."""
    _require_cuda(vals, logp, db)
    _require_cuda_any(rows, cols, dz_bf16)
    m = logp.shape[0]
    assert rows.dtype == torch.int64 and cols.dtype == torch.int64 and rows.numel() == m and cols.numel() == m
    assert rows.is_contiguous() and cols.is_contiguous() and vals.is_contiguous() and vals.numel() == m
    assert logp.is_contiguous() and dz_bf16.dtype == torch.bfloat16 and dz_bf16.shape[0] == m and dz_bf16.stride(1) == 1
    _check(load().hcgs_logsoftmax_bwd_fused_onehot(_ptr(rows), _ptr(cols), _ptr(vals), _ptr(logp), m, logp.shape[1],
                                                   _ptr(dz_bf16), dz_bf16.stride(0), _ptr(db), _stream()),
           'hcgs_logsoftmax_bwd_fused_onehot')


def nll_err_rows(logp, labels):
    """(-logp[i, label_i], first-argmax(logp[i]) != label_i) per row, one pass; two fp32 CUDA vectors."""
    _require_cuda(logp)
    _require_cuda_any(labels)
    assert logp.dim() == 2 and logp.stride(1) == 1 and labels.dtype == torch.int64 and labels.is_contiguous()
    m = logp.shape[0]
    assert labels.numel() == m
    loss_rows = torch.empty(m, device=logp.device, dtype=torch.float32)
    err_rows = torch.empty(m, device=logp.device, dtype=torch.float32)
    _check(load().hcgs_nll_err_rows(_ptr(logp), logp.stride(0), m, logp.shape[1], _ptr(labels), _ptr(loss_rows),
                                    _ptr(err_rows), _stream()), 'hcgs_nll_err_rows')
    return loss_rows, err_rows


def nll_fwd_bwd(logp, labels, ldlab, scale, dz, stats):
    _require_cuda(logp, labels, dz, stats)
    assert logp.is_contiguous()
    _check(load().hcgs_nll_fwd_bwd(_ptr(logp), logp.shape[0], logp.shape[1], _ptr(labels), int(ldlab), float(scale),
                                   _ptr(dz), _ptr(stats), _stream()), 'hcgs_nll_fwd_bwd')


OPT_KINDS = {'sgd': 0, 'rmsprop': 1, 'adam': 2}


def opt_step(kind, param, grad, state1, state2, hp, grad_scale=1.0):
    _require_cuda(param, grad, state1, state2)
    assert param.is_contiguous() and grad.is_contiguous() and len(hp) == 8
    arr = (c_float * 8)(*[float(v) for v in hp])
    _check(load().hcgs_opt_step(OPT_KINDS[kind], _ptr(param), _ptr(grad), _ptr(state1), _ptr(state2), param.numel(),
                                arr, float(grad_scale), _stream()), 'hcgs_opt_step')


def opt_step_multi(kind, params, grads, state1, state2, hp, grad_scale=1.0):
    """One fused optimizer update for a list of fp32 CUDA tensors (one launch per 48 tensors).  state1 / state2: lists of
    tensors or None (kinds that do not use them)."""
    _require_cuda(*params, *grads, *(state1 or ()), *(state2 or ()))
    n = len(params)
    assert len(grads) == n and len(hp) == 8 and all(t.is_contiguous() for t in list(params) + list(grads))

    def table(ts):
        return (c_void_p * n)(*[t.data_ptr() for t in ts]) if ts else None
    arr = (ctypes.c_double * 8)(*[float(v) for v in hp])
    numels = (c_int64 * n)(*[int(p.numel()) for p in params])
    _check(load().hcgs_opt_step_multi(OPT_KINDS[kind], n, table(params), table(grads), table(state1), table(state2),
                                      numels, arr, float(grad_scale), _stream()), 'hcgs_opt_step_multi')


def assemble_batch(chunk, beg, length, left, max_len, out=None):
    """out[t, k, :] = chunk[beg[k] + t - left[k], :] inside sentence k, zeros elsewhere (core.py:183-200 in one launch).
    chunk: [rows, dim] fp32 CUDA; beg int64, length / left int32 CUDA vectors of `batch` entries."""
    _require_cuda(chunk)
    _require_cuda_any(beg, length, left)
    assert chunk.dim() == 2 and chunk.is_contiguous() and chunk.dtype == torch.float32
    assert beg.dtype == torch.int64 and length.dtype == torch.int32 and left.dtype == torch.int32
    batch = beg.numel()
    assert length.numel() == batch and left.numel() == batch and batch >= 1
    assert beg.is_contiguous() and length.is_contiguous() and left.is_contiguous()
    dim = chunk.shape[1]
    if out is None:
        out = torch.empty(int(max_len), batch, dim, device=chunk.device, dtype=torch.float32)
    assert out.is_contiguous() and tuple(out.shape) == (int(max_len), batch, dim)
    _check(load().hcgs_assemble_batch(_ptr(chunk), chunk.shape[0], dim, _ptr(beg), _ptr(length), _ptr(left), batch,
                                      int(max_len), _ptr(out), _stream()), 'hcgs_assemble_batch')
    return out


def unpad_rows(src, row_off, length, t_max, out):
    """out[row_off[k] + t, :] = src[t * batch + k, :] for t < length[k]: per-utterance matrices, padding dropped.
    src [t_max * batch, cols] fp32 CUDA (row stride = src.stride(0)); out [>= sum(length), cols] contiguous."""
    _require_cuda(src, out)
    _require_cuda_any(row_off, length)
    assert row_off.dtype == torch.int64 and length.dtype == torch.int32 and row_off.is_contiguous() and length.is_contiguous()
    batch = row_off.numel()
    assert length.numel() == batch and src.dim() == 2 and src.stride(1) == 1 and src.shape[0] == int(t_max) * batch
    cols = src.shape[1]
    assert out.is_contiguous() and out.dim() == 2 and out.shape[1] == cols
    _check(load().hcgs_unpad_rows(_ptr(src), src.stride(0), cols, _ptr(row_off), _ptr(length), batch, int(t_max),
                                  _ptr(out), _stream()), 'hcgs_unpad_rows')
    return out


def window_abs_mean(w, windows):
    """mean |w| over each window (r0, r1, c0, c1) of the int32 CUDA tensor `windows` [n, 4]; returns a CUDA float vector."""
    _require_cuda(w)
    _require_cuda_any(windows)
    assert w.dim() == 2 and w.stride(1) == 1 and windows.dtype == torch.int32 and windows.is_contiguous()
    n = windows.shape[0]
    means = torch.empty(n, device=w.device, dtype=torch.float32)
    _check(load().hcgs_window_abs_mean(_ptr(w), w.stride(0), _ptr(windows), n, _ptr(means), _stream()),
           'hcgs_window_abs_mean')
    return means


def fill_rects(out, rects, value=1.0):
    _require_cuda(out)
    _require_cuda_any(rects)
    assert out.dim() == 2 and out.stride(1) == 1 and rects.dtype == torch.int32 and rects.is_contiguous()
    _check(load().hcgs_fill_rects(_ptr(out), out.stride(0), _ptr(rects), rects.shape[0], float(value), _stream()),
           'hcgs_fill_rects')
    return out

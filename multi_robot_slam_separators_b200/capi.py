"""ctypes binding of include/sepb200.h (the C-ABI shared library libsepb200.so).

The library is the product; this module only marshals numpy buffers.  There is no CPU
fallback: if the library is missing or no sm_100 device is present, calls raise.
"""
from __future__ import annotations

import ctypes as C
import weakref
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsepb200.so")

_lib = None


class SepbError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SepbError(f"{LIB_PATH} not built: run `python -m multi_robot_slam_separators_b200.build` "
                            "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _declare(_lib)
    return _lib


class MatchParams(C.Structure):
    _fields_ = [("nndr", C.c_float), ("nndr_strict", C.c_int32)]


class PnpParams(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("iterations", C.c_int32), ("min_inliers", C.c_int32), ("reproj_error", C.c_double),
                ("confidence", C.c_double), ("guess_rvec", C.c_double * 3), ("guess_tvec", C.c_double * 3),
                ("local_R", C.c_double * 9), ("local_t", C.c_double * 3),
                ("flags", C.c_int32), ("refine_iterations", C.c_int32), ("refine_sigma", C.c_double),
                ("minimal_model", C.c_int32), ("reserved_", C.c_int32)]


class GuessParams(C.Structure):
    _fields_ = [("image_width", C.c_int32), ("image_height", C.c_int32), ("win_size", C.c_float)]


class BaParams(C.Structure):
    _fields_ = [("type", C.c_int32), ("iterations", C.c_int32), ("pixel_variance", C.c_double),
                ("robust_kernel_delta", C.c_double), ("baseline", C.c_double), ("variance_ignored", C.c_int32),
                ("reserved_", C.c_int32)]


class VerifyOut(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("status", "n_matches", "n_inliers", "pose", "transform", "match_from",
                                          "match_to", "inlier_mask", "best_hyp", "iters_run", "covariance")]


PNP_ITERATIVE, PNP_P3P, PNP_AP3P = 0, 2, 5      # sepb_pnp_params.flags (cv::SolvePnPMethod values, Vis/PnPFlags)


def pnp_params(fx, fy, cx, cy, iterations=300, min_inliers=5, reproj_error=2.0, confidence=0.99, guess=None,
               flags=PNP_ITERATIVE, refine_iterations=0, refine_sigma=3.0, minimal_model=False):
    p = PnpParams()
    lib().sepb_pnp_params_init(C.byref(p), fx, fy, cx, cy)
    p.iterations, p.min_inliers, p.reproj_error, p.confidence = iterations, min_inliers, reproj_error, confidence
    p.flags, p.refine_iterations, p.refine_sigma, p.minimal_model = flags, refine_iterations, refine_sigma, int(minimal_model)
    if guess is not None:
        g = np.ascontiguousarray(np.asarray(guess, dtype=np.float64)[:3, :4])
        lib().sepb_pnp_set_guess(C.byref(p), _ptr(g))
    return p


def ba_params(baseline, iterations=20, pixel_variance=1.0, robust_kernel_delta=8.0, variance_ignored=False):
    """sepb_ba_params: Vis/BundleAdjustment = 1 with RTAB-Map's defaults; `baseline` = stereo baseline * fy / fx [m]."""
    b = BaParams()
    lib().sepb_ba_params_init(C.byref(b), baseline)
    b.iterations, b.pixel_variance, b.robust_kernel_delta = iterations, pixel_variance, robust_kernel_delta
    b.variance_ignored = int(variance_ignored)
    return b


def _declare(L):
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    L.sepb_ctx_create.argtypes = [i32, C.POINTER(vp)]
    L.sepb_ctx_create.restype = i32
    L.sepb_ctx_create_priority.argtypes = [i32, C.POINTER(vp)]
    L.sepb_ctx_create_priority.restype = i32
    L.sepb_ctx_destroy.argtypes = [vp]
    L.sepb_ctx_destroy.restype = None
    L.sepb_last_error.argtypes = [vp]
    L.sepb_last_error.restype = C.c_char_p
    L.sepb_ctx_stream.argtypes = [vp]
    L.sepb_ctx_stream.restype = vp
    L.sepb_ctx_sync.argtypes = [vp]
    L.sepb_ctx_sync.restype = i32
    L.sepb_ctx_launch_count.argtypes = [vp]
    L.sepb_ctx_launch_count.restype = i64
    L.sepb_ctx_set_option.argtypes = [vp, i32, i32]
    L.sepb_ctx_set_option.restype = i32
    L.sepb_ctx_get_option.argtypes = [vp, i32, C.POINTER(C.c_int)]
    L.sepb_ctx_get_option.restype = i32
    L.sepb_frames_create.argtypes = [vp, i32, i32, C.POINTER(vp)]
    L.sepb_frames_create.restype = i32
    L.sepb_frames_destroy.argtypes = [vp]
    L.sepb_frames_destroy.restype = None
    L.sepb_frames_size.argtypes = [vp]
    L.sepb_frames_size.restype = i32
    L.sepb_frames_fetched_bytes.argtypes = [vp, vp]
    L.sepb_frames_fetched_bytes.restype = C.c_int64
    for name in ("sepb_frames_upload", "sepb_frames_upload_async", "sepb_frames_bind_host_async"):
        f = getattr(L, name)
        f.argtypes = [vp, vp, i32, i32, vp, vp, vp, vp, vp]
        f.restype = i32
    L.sepb_frames_wait.argtypes = [vp, vp]
    L.sepb_frames_wait.restype = i32
    L.sepb_measure_peak.argtypes = [vp, i32, i32, C.POINTER(C.c_double)]
    L.sepb_measure_peak.restype = i32
    L.sepb_pnp_params_init.argtypes = [C.POINTER(PnpParams), C.c_double, C.c_double, C.c_double, C.c_double]
    L.sepb_pnp_params_init.restype = None
    L.sepb_pnp_set_guess.argtypes = [C.POINTER(PnpParams), vp]
    L.sepb_pnp_set_guess.restype = None
    L.sepb_pnp_ransac_dev.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, vp, C.POINTER(PnpParams), vp, i32, vp,
                                      vp, vp, vp, vp, vp, i32, vp, vp]
    L.sepb_pnp_ransac_dev.restype = i32
    L.sepb_db_create.argtypes = [vp, i32, i32, C.POINTER(vp)]
    L.sepb_db_create.restype = i32
    L.sepb_db_destroy.argtypes = [vp]
    L.sepb_db_destroy.restype = None
    L.sepb_db_size.argtypes = [vp]
    L.sepb_db_size.restype = i32
    L.sepb_db_clear.argtypes = [vp]
    L.sepb_db_clear.restype = i32
    L.sepb_db_append.argtypes = [vp, vp, vp, i32]
    L.sepb_db_append.restype = i32
    L.sepb_db_set_base.argtypes = [vp, i64]
    L.sepb_db_set_base.restype = i32
    L.sepb_db_set_col_mask.argtypes = [vp, vp, vp]
    L.sepb_db_set_col_mask.restype = i32
    L.sepb_netvlad_topk.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp, i32, vp, vp]
    L.sepb_netvlad_topk.restype = i32
    L.sepb_netvlad_topk_dev.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp, vp, vp, vp]
    L.sepb_netvlad_topk_dev.restype = i32
    L.sepb_netvlad_topk_exact_static_dev.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, vp, vp]
    L.sepb_netvlad_topk_exact_static_dev.restype = i32
    L.sepb_netvlad_filter_debug.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, vp]
    L.sepb_netvlad_filter_debug.restype = i32
    L.sepb_topk_merge_dev.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp]
    L.sepb_topk_merge_dev.restype = i32
    L.sepb_select_matches.argtypes = [vp, vp, i32, C.c_double, i32, vp]
    L.sepb_select_matches.restype = i32
    for name in ("sepb_verify_pairs", "sepb_verify_pairs_dev", "sepb_verify_pairs_async"):
        f = getattr(L, name)
        f.argtypes = [vp, vp, vp, vp, i32, C.POINTER(MatchParams), C.POINTER(PnpParams), vp, i32,
                      C.POINTER(VerifyOut)]
        f.restype = i32
    L.sepb_guess_params_init.argtypes = [C.POINTER(GuessParams), i32, i32]
    L.sepb_guess_params_init.restype = None
    L.sepb_guess_match_pairs.argtypes = [vp, vp, vp, vp, i32, vp, C.POINTER(PnpParams), C.POINTER(GuessParams), C.c_float,
                                         vp, vp, vp, vp]
    L.sepb_guess_match_pairs.restype = i32
    for name in ("sepb_estimate_transformations", "sepb_estimate_transformations_dev"):
        f = getattr(L, name)
        f.argtypes = [vp, vp, vp, vp, i32, C.POINTER(MatchParams), C.POINTER(PnpParams), C.POINTER(GuessParams),
                      C.POINTER(VerifyOut), vp]
        f.restype = i32
    for name in ("sepb_match_pairs", "sepb_match_pairs_dev"):
        f = getattr(L, name)
        f.argtypes = [vp, vp, vp, vp, i32, C.POINTER(MatchParams), vp, vp, vp, vp, vp, vp]
        f.restype = i32
    L.sepb_ba_params_init.argtypes = [C.POINTER(BaParams), C.c_double]
    L.sepb_ba_params_init.restype = None
    L.sepb_ctx_set_bundle_adjustment.argtypes = [vp, C.POINTER(BaParams)]
    L.sepb_ctx_set_bundle_adjustment.restype = i32
    for name in ("sepb_bundle_adjust_pairs", "sepb_bundle_adjust_dev"):
        f = getattr(L, name)
        f.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, vp, C.POINTER(PnpParams), C.POINTER(BaParams), vp, vp, vp, vp, vp,
                      i32, vp]
        f.restype = i32
    L.sepb_group_create.argtypes = [i32, vp, C.POINTER(vp)]
    L.sepb_group_create.restype = i32
    L.sepb_group_unique_id.argtypes = [vp]
    L.sepb_group_unique_id.restype = i32
    L.sepb_group_create_rank.argtypes = [vp, vp, i32, i32, C.POINTER(vp)]
    L.sepb_group_create_rank.restype = i32
    L.sepb_group_destroy.argtypes = [vp]
    L.sepb_group_destroy.restype = None
    L.sepb_group_local_size.argtypes = [vp]
    L.sepb_group_local_size.restype = i32
    L.sepb_group_nranks.argtypes = [vp]
    L.sepb_group_nranks.restype = i32
    L.sepb_group_ctx.argtypes = [vp, i32]
    L.sepb_group_ctx.restype = vp
    L.sepb_group_last_error.argtypes = [vp]
    L.sepb_group_last_error.restype = C.c_char_p
    L.sepb_group_nccl_version.argtypes = []
    L.sepb_group_nccl_version.restype = i32
    L.sepb_group_last_gather_bytes.argtypes = [vp]
    L.sepb_group_last_gather_bytes.restype = i64
    L.sepb_group_netvlad_topk_dev.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp]
    L.sepb_group_netvlad_topk_dev.restype = i32
    L.sepb_group_netvlad_topk.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp]
    L.sepb_group_netvlad_topk.restype = i32


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    assert a.flags["C_CONTIGUOUS"]
    return C.c_void_p(a.ctypes.data)


OPT_MATCH_KERNEL = 1
OPT_PNP_SCHEDULER = 2
OPT_MATCH_TILE = 3
OPT_MATCH_SCHEDULER = 4


class Context:
    def __init__(self, device: int = 0, _borrowed=None, high_priority: bool = False):
        """_borrowed: handle of a context owned by a sepb_group (close() then only closes the children).
        high_priority: streams at the device's highest priority (short searches that must overtake a long batch)."""
        self._owned = _borrowed is None
        if _borrowed is not None:
            self._h = C.c_void_p(_borrowed)
        else:
            self._h = C.c_void_p()
            rc = (lib().sepb_ctx_create_priority if high_priority else lib().sepb_ctx_create)(device, C.byref(self._h))
            if rc != 0:
                raise SepbError(f"sepb_ctx_create(device={device}) failed with code {rc} (CUDA device required)")
        # stores / databases created on this context: they hold a raw pointer to it, so close() destroys them first
        self._children = weakref.WeakSet()

    def check(self, rc: int):
        if rc != 0:
            raise SepbError(f"sepb error {rc}: {lib().sepb_last_error(self._h).decode()}")

    @property
    def stream(self) -> int:
        return lib().sepb_ctx_stream(self._h) or 0

    def sync(self):
        self.check(lib().sepb_ctx_sync(self._h))

    @property
    def launch_count(self) -> int:
        return int(lib().sepb_ctx_launch_count(self._h))

    def measure_peak(self, kind: str, reps: int = 5) -> float:
        """'popc' -> POPC/s, 'dfma' -> FP64 FMA/s, 'tmem' / 'tmem_pack16' -> TMEM bytes/s read by tcgen05.ld
        (measured microbenchmarks); 'mma_i8_n128' | 'mma_i8_n256' | 'mma_bf16_n256' | 'mma_i8_n128_sts' -> tensor-pipe
        ops/s of a bare tcgen05.mma issue loop (M128, SWIZZLE_128B shared-memory operands; _sts: with producer-like
        shared-memory stores running beside it); 'mma2_i8_m256n256' | 'mma2_bf16_m256n256' -> the same loop issued by CTA
        pairs (tcgen05.mma.cta_group::2, M256 x N256)."""
        v = C.c_double()
        self.check(lib().sepb_measure_peak(self._h, {"popc": 0, "dfma": 1, "tmem": 2, "tmem_pack16": 3, "mma_i8_n128": 4, "mma_i8_n256": 5,
                                                         "mma_bf16_n256": 6, "mma_i8_n128_sts": 7, "mma2_i8_m256n256": 8,
                                                         "mma2_bf16_m256n256": 9}[kind], reps, C.byref(v)))
        return v.value

    def set_match_kernel(self, which: str):
        """'auto' | 'popc' (INT-pipe kernel) | 'tensor' (tcgen05 int8 GEMM kernel); identical results."""
        self.check(lib().sepb_ctx_set_option(self._h, OPT_MATCH_KERNEL, {"auto": -1, "popc": 0, "tensor": 1}[which]))

    def set_match_tile(self, width: int):
        """Accumulator tile of the tensor matcher: 0 = widest that fits (256 columns up to ~1400 features), 128 = narrow."""
        self.check(lib().sepb_ctx_set_option(self._h, OPT_MATCH_TILE, int(width)))

    def set_match_scheduler(self, which: str):
        """Tensor matcher: 'persistent' (default: one CTA per SM runs through its pairs, tails overlap the next pair's
        first tiles) | 'per_pair' (one CTA per pair); identical results."""
        self.check(lib().sepb_ctx_set_option(self._h, OPT_MATCH_SCHEDULER, {"persistent": 0, "per_pair": 1}[which]))

    def set_bundle_adjustment(self, ba: "BaParams | None"):
        """Vis/BundleAdjustment for verify_pairs / estimate_transformations on this context (None: off, the default)."""
        self.check(lib().sepb_ctx_set_bundle_adjustment(self._h, C.byref(ba) if ba is not None else None))

    def set_pnp_scheduler(self, which: str):
        """'persistent' (default: persistent CTAs, device work list, idle warps compute chunks of running pairs) |
        'per_pair' (one one-warp CTA per pair); identical results."""
        self.check(lib().sepb_ctx_set_option(self._h, OPT_PNP_SCHEDULER, {"persistent": 0, "per_pair": 1}[which]))

    def close(self):
        if self._h:
            for child in list(getattr(self, "_children", ())):
                child.close()
            if getattr(self, "_owned", True):
                lib().sepb_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Frames:
    """Device-resident feature store of one robot (sepb_frames)."""

    def __init__(self, ctx: Context, capacity: int, max_feat: int):
        self.ctx, self.capacity, self.max_feat = ctx, capacity, max_feat
        self._h = C.c_void_p()
        ctx.check(lib().sepb_frames_create(ctx._h, capacity, max_feat, C.byref(self._h)))
        ctx._children.add(self)

    def upload(self, first, desc, kp2d, kp3d, count, octave=None, async_=False, lazy=False):
        """async_=True: enqueue on the copy stream and return (arrays must stay alive, ideally pinned, until
        wait() or the next blocking call that uses the store).
        lazy=True (sepb_frames_bind_host_async; implies async_): only descriptors + counts are copied, kp2d / kp3d /
        octave must be PINNED numpy arrays of exactly the store's layout and are read in place by the device for the
        frames whose pairs reach PnP; they must stay alive and unchanged until the next upload of these slots."""
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        n = desc.shape[0]
        assert desc.shape == (n, self.max_feat, 32)
        kp2d = np.ascontiguousarray(kp2d, dtype=np.float32).reshape(n, self.max_feat, 2)
        kp3d = np.ascontiguousarray(kp3d, dtype=np.float32).reshape(n, self.max_feat, 3)
        count = np.ascontiguousarray(count, dtype=np.int32).reshape(n)
        if octave is not None:
            octave = np.ascontiguousarray(octave, dtype=np.int32).reshape(n, self.max_feat)
        fn = lib().sepb_frames_bind_host_async if lazy else (lib().sepb_frames_upload_async if async_ else lib().sepb_frames_upload)
        self.ctx.check(fn(self.ctx._h, self._h, first, n, _ptr(desc), _ptr(kp2d), _ptr(kp3d), _ptr(octave), _ptr(count)))
        if async_ or lazy:
            self._pending = (desc, kp2d, kp3d, count, octave)     # keep the host arrays alive
        if lazy:
            self._bound = (kp2d, kp3d, octave)                    # read in place by the device until the next upload

    @property
    def fetched_bytes(self) -> int:
        """Bytes the device pulled from the bound host arrays since the last lazy upload (synchronises)."""
        return int(lib().sepb_frames_fetched_bytes(self.ctx._h, self._h))

    def wait(self):
        self.ctx.check(lib().sepb_frames_wait(self.ctx._h, self._h))
        self._pending = None

    @property
    def size(self):
        return lib().sepb_frames_size(self._h)

    def close(self):
        if self._h:
            lib().sepb_frames_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def frames_from_robot(ctx: Context, robot) -> Frames:
    """Upload a synth.RobotFrames."""
    n, F = robot.desc.shape[0], robot.desc.shape[1]
    f = Frames(ctx, n, F)
    f.upload(0, robot.desc, robot.kp2d, robot.kp3d, robot.count, robot.octave)
    return f


def match_pairs(ctx: Context, frames_from: Frames, frames_to: Frames, pairs, nndr=0.8, strict=False, want_knn=True):
    """Host-buffer K2 call.  pairs: [P,2] (from_frame, to_frame).  Returns dict of numpy arrays."""
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    P = pairs.shape[0]
    Ft, Ff = frames_to.max_feat, frames_from.max_feat
    knn_idx = np.full((P, Ft, 2), -1, np.int32) if want_knn else None
    knn_dist = np.full((P, Ft, 2), -1, np.int32) if want_knn else None
    n_matches = np.zeros(P, np.int32)
    n_unique = np.zeros(P, np.int32)
    mf = np.zeros((P, Ff), np.uint16)
    mt = np.zeros((P, Ff), np.uint16)
    prm = MatchParams(nndr, 1 if strict else 0)
    ctx.check(lib().sepb_match_pairs(ctx._h, frames_from._h, frames_to._h, _ptr(pairs), P, C.byref(prm),
                                     _ptr(knn_idx), _ptr(knn_dist), _ptr(n_matches), _ptr(mf), _ptr(mt),
                                     _ptr(n_unique)))
    return dict(knn_idx=knn_idx, knn_dist=knn_dist, n_matches=n_matches, match_from=mf, match_to=mt,
                n_unique_to=n_unique)


def verify_buffers(P, F, want_lists=True, alloc=np.zeros):
    """Result arrays of verify_pairs; `alloc(shape, dtype)` lets a caller hand in pinned memory."""
    res = dict(status=alloc(P, np.int32), n_matches=alloc(P, np.int32), n_inliers=alloc(P, np.int32),
               pose=alloc((P, 6), np.float64), transform=alloc((P, 12), np.float32),
               best_hyp=alloc(P, np.int32), iters_run=alloc(P, np.int32))
    if want_lists:
        res.update(match_from=alloc((P, F), np.uint16), match_to=alloc((P, F), np.uint16),
                   inlier_mask=alloc((P, (F + 31) // 32), np.uint32), covariance=alloc((P, 36), np.float64))
    return res


def verify_pairs(ctx: Context, frames_from: Frames, frames_to: Frames, pairs, pnp: PnpParams, nndr=0.8, strict=False,
                 custom_samples=None, want_lists=True, out=None, async_=False):
    """Host-buffer fused K2 -> K3 call (estimate_transformation pass 1 for a batch of pairs).
    want_lists=False skips the D2H of match lists / inlier masks (only what EstTransform.srv returns).
    async_=True (sepb_verify_pairs_async): returns once enqueued; `pairs` and `out` must be pinned and the results
    are valid after ctx.sync()."""
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    P, F = pairs.shape[0], frames_from.max_feat
    res = out if out is not None else verify_buffers(P, F, want_lists)
    out = VerifyOut(**{k: v.ctypes.data for k, v in res.items()})
    cs, ch = None, 0
    if custom_samples is not None:
        cs = np.ascontiguousarray(custom_samples, dtype=np.int32)
        assert cs.ndim == 3 and cs.shape[0] == P and cs.shape[2] == (5 if pnp.flags == PNP_ITERATIVE else 4)
        ch = cs.shape[1]
    mp = MatchParams(nndr, 1 if strict else 0)
    fn = lib().sepb_verify_pairs_async if async_ else lib().sepb_verify_pairs
    ctx.check(fn(ctx._h, frames_from._h, frames_to._h, _ptr(pairs), P, C.byref(mp), C.byref(pnp), _ptr(cs), ch,
                 C.byref(out)))
    return res


def bundle_adjust_pairs(ctx: Context, frames_from: Frames, frames_to: Frames, pairs, res, pnp: PnpParams, ba: BaParams,
                        use_covariance=True):
    """sepb_bundle_adjust_pairs on a COPY of the verify_pairs result `res` (needs its match lists, masks and, with
    use_covariance, its covariance): returns the updated status / n_inliers / pose / transform / inlier_mask + stats."""
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    P = pairs.shape[0]
    out = {k: np.ascontiguousarray(res[k]).copy() for k in ("status", "n_inliers", "pose", "transform", "inlier_mask")}
    out["stats"] = np.zeros((P, 4), np.float64)
    cov = np.ascontiguousarray(res["covariance"], dtype=np.float64) if use_covariance else None
    ctx.check(lib().sepb_bundle_adjust_pairs(
        ctx._h, frames_from._h, frames_to._h, _ptr(pairs), P, _ptr(np.ascontiguousarray(res["n_matches"], dtype=np.int32)),
        _ptr(np.ascontiguousarray(res["match_from"])), _ptr(np.ascontiguousarray(res["match_to"])), _ptr(cov),
        C.byref(pnp), C.byref(ba), _ptr(out["status"]), _ptr(out["n_inliers"]), _ptr(out["pose"]), _ptr(out["transform"]),
        _ptr(out["inlier_mask"]), out["inlier_mask"].shape[1], _ptr(out["stats"])))
    return out


def guess_params(width, height, win_size=20.0):
    g = GuessParams()
    lib().sepb_guess_params_init(C.byref(g), int(width), int(height))
    g.win_size = float(win_size)
    return g


def guess_match_pairs(ctx, frames_from, frames_to, pairs, transforms, pnp: PnpParams, gprm: GuessParams, nndr=0.8):
    """K2G on host buffers: pass-2 correspondence lists for explicit first-pass transforms [P,12] f32."""
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    P, F = pairs.shape[0], frames_from.max_feat
    T = np.ascontiguousarray(transforms, dtype=np.float32).reshape(P, 12)
    nm = np.zeros(P, np.int32)
    mf, mt = np.zeros((P, F), np.uint16), np.zeros((P, F), np.uint16)
    g = np.zeros((P, 6), np.float64)
    ctx.check(lib().sepb_guess_match_pairs(ctx._h, frames_from._h, frames_to._h, _ptr(pairs), P, _ptr(T), C.byref(pnp),
                                           C.byref(gprm), float(nndr), _ptr(nm), _ptr(mf), _ptr(mt), _ptr(g)))
    return dict(n_matches=nm, match_from=mf, match_to=mt, guess=g)


def estimate_transformations(ctx, frames_from, frames_to, pairs, pnp: PnpParams, gprm: GuessParams, nndr=0.8,
                             want_lists=True):
    """The batched estimate_transformation service (both calls of stereoCamGeometricTools.cpp:141-164)."""
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    P, F = pairs.shape[0], frames_from.max_feat
    res = verify_buffers(P, F, want_lists)
    out = VerifyOut(**{k: v.ctypes.data for k, v in res.items()})
    st1 = np.zeros(P, np.int32)
    mp = MatchParams(nndr, 0)
    ctx.check(lib().sepb_estimate_transformations(ctx._h, frames_from._h, frames_to._h, _ptr(pairs), P, C.byref(mp),
                                                  C.byref(pnp), C.byref(gprm), C.byref(out), _ptr(st1)))
    res["status_pass1"] = st1
    return res


def verify_pairs_light(ctx, frames_from, frames_to, pairs, pnp, nndr=0.8):
    return verify_pairs(ctx, frames_from, frames_to, pairs, pnp, nndr, want_lists=False)


def inlier_indices(mask_row, n_matches):
    """bitmask over match-list positions -> ascending index array"""
    bits = np.unpackbits(mask_row.view(np.uint8), bitorder="little")[:n_matches]
    return np.nonzero(bits)[0].astype(np.int32)


NN_EXACT, NN_TENSOR = 0, 1


class NetvladDB:
    """Device-resident, append-only NetVLAD descriptor database (received_descriptors, data_handler.py:300-301)."""

    def __init__(self, ctx: Context, dim: int, capacity: int, base: int = 0):
        self.ctx, self.dim, self.capacity = ctx, dim, capacity
        self._h = C.c_void_p()
        ctx.check(lib().sepb_db_create(ctx._h, dim, capacity, C.byref(self._h)))
        ctx._children.add(self)
        if base:
            ctx.check(lib().sepb_db_set_base(self._h, base))

    def append(self, rows):
        rows = np.ascontiguousarray(rows, dtype=np.float64).reshape(-1, self.dim)
        self.ctx.check(lib().sepb_db_append(self.ctx._h, self._h, _ptr(rows), rows.shape[0]))

    @property
    def size(self):
        return lib().sepb_db_size(self._h)

    def clear(self):
        self.ctx.check(lib().sepb_db_clear(self._h))

    def set_col_mask(self, used_indices):
        """other_kf_already_used (list of database row indices) -> byte mask."""
        if used_indices is None or len(used_indices) == 0:
            self.ctx.check(lib().sepb_db_set_col_mask(self.ctx._h, self._h, None))
            return
        m = np.zeros(self.size, np.uint8)
        m[np.asarray(used_indices, dtype=np.int64)] = 1
        self.ctx.check(lib().sepb_db_set_col_mask(self.ctx._h, self._h, _ptr(m)))

    def topk(self, queries, k=1, mode=NN_EXACT, local_used=(), pairs_ignored=()):
        q = np.ascontiguousarray(queries, dtype=np.float64).reshape(-1, self.dim)
        nq = q.shape[0]
        rm = None
        if len(local_used):
            rm = np.zeros(nq, np.uint8)
            rm[np.asarray(local_used, dtype=np.int64)] = 1
        ign = np.ascontiguousarray(pairs_ignored, dtype=np.int32).reshape(-1, 2) if len(pairs_ignored) else None
        idx = np.full((nq, k), -1, np.int32)
        dist = np.full((nq, k), np.inf, np.float64)
        self.ctx.check(lib().sepb_netvlad_topk(self.ctx._h, self._h, _ptr(q), nq, k, mode, _ptr(rm), _ptr(ign),
                                               0 if ign is None else ign.shape[0], _ptr(idx), _ptr(dist)))
        return idx, dist

    def topk_exact_static_dev(self, d_queries: int, nq: int, k: int, d_out_idx: int, d_out_dist: int, d_row_mask: int = 0,
                              d_ign_start: int = 0, d_ign_col: int = 0):
        """sepb_netvlad_topk_exact_static_dev: the exact search on DEVICE pointers (integers) with the partial lists in
        the library's fixed 8 MiB device array -- no allocation inside the call.  Asynchronous on the context's stream;
        overlapping calls on one device race on that array."""
        p = lambda a: C.c_void_p(a) if a else None  # noqa: E731
        self.ctx.check(lib().sepb_netvlad_topk_exact_static_dev(self.ctx._h, self._h, p(d_queries), nq, k, p(d_row_mask),
                                                                p(d_ign_start), p(d_ign_col), p(d_out_idx), p(d_out_dist)))

    def filter_debug(self, queries, k=5, want_v=True):
        """K1 tensor-core filter with its raw values exposed (tests / profiling):
        (v [nq, size] f32 or None, idx, dist, stats[4])."""
        q = np.ascontiguousarray(queries, dtype=np.float64).reshape(-1, self.dim)
        nq = q.shape[0]
        v = np.empty((nq, self.size), np.float32) if want_v else None
        idx = np.full((nq, k), -1, np.int32)
        dist = np.full((nq, k), np.inf, np.float64)
        stats = np.zeros(4, np.int32)
        self.ctx.check(lib().sepb_netvlad_filter_debug(self.ctx._h, self._h, _ptr(q), nq, k, _ptr(v), _ptr(idx),
                                                       _ptr(dist), _ptr(stats)))
        return v, idx, dist, stats

    def close(self):
        if self._h:
            lib().sepb_db_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def select_matches(row_min_idx, row_min_dist, netvlad_distance=0.13, max_matches_nb=20):
    """data_handler.py:191-205 (host helper inside the C ABI)."""
    i = np.ascontiguousarray(row_min_idx, dtype=np.int32).ravel()
    d = np.ascontiguousarray(row_min_dist, dtype=np.float64).ravel()
    out = np.zeros((max(max_matches_nb, 1), 2), np.int32)
    n = lib().sepb_select_matches(_ptr(i), _ptr(d), len(i), float(netvlad_distance), int(max_matches_nb), _ptr(out))
    if n < 0:
        raise SepbError(f"sepb_select_matches failed: {n}")
    return [(int(a), int(b)) for a, b in out[:n]]


class Group:
    """sepb_group: the devices of one box and their NCCL communicator (include/sepb200.h, multi-GPU section).

    Group(ndev=N)                       one process drives N devices (ncclCommInitAll);
    Group(ctx=c, rank=r, nranks=N, unique_id=b)   one process per device (ncclCommInitRank); `unique_id` = 128 bytes made
                                        by Group.unique_id() on rank 0 and handed to every rank by the launcher."""

    def __init__(self, ndev: int | None = None, devices=None, ctx: Context | None = None, rank: int = 0, nranks: int = 1,
                 unique_id: bytes | None = None):
        self._h = C.c_void_p()
        if ctx is None:
            ndev = len(devices) if devices is not None else int(ndev or 1)
            dv = np.ascontiguousarray(devices, dtype=np.int32) if devices is not None else None
            rc = lib().sepb_group_create(ndev, _ptr(dv), C.byref(self._h))
            if rc != 0:
                raise SepbError(f"sepb_group_create({ndev}) failed with code {rc} (-3: NCCL not available, -4: NCCL error)")
            self.ctxs = [Context(_borrowed=lib().sepb_group_ctx(self._h, i)) for i in range(ndev)]
        else:
            uid = np.frombuffer(unique_id if unique_id is not None else bytes(128), dtype=np.uint8).copy()
            assert uid.size == 128
            rc = lib().sepb_group_create_rank(ctx._h, _ptr(uid), int(rank), int(nranks), C.byref(self._h))
            if rc != 0:
                raise SepbError(f"sepb_group_create_rank(rank={rank}, nranks={nranks}) failed with code {rc}")
            self.ctxs = [ctx]
        self.nranks = lib().sepb_group_nranks(self._h)

    @staticmethod
    def unique_id() -> bytes:
        uid = np.zeros(128, np.uint8)
        rc = lib().sepb_group_unique_id(_ptr(uid))
        if rc != 0:
            raise SepbError(f"sepb_group_unique_id failed with code {rc} (NCCL not available?)")
        return uid.tobytes()

    @staticmethod
    def nccl_version() -> int:
        return int(lib().sepb_group_nccl_version())

    def check(self, rc: int):
        if rc != 0:
            raise SepbError(f"sepb group error {rc}: {lib().sepb_group_last_error(self._h).decode()}")

    @property
    def last_gather_bytes(self) -> int:
        return int(lib().sepb_group_last_gather_bytes(self._h))

    def _handles(self, dbs):
        assert len(dbs) == len(self.ctxs)
        return (C.c_void_p * len(dbs))(*[d._h for d in dbs])

    def netvlad_topk(self, dbs, queries, k=5, mode=1):
        """Host form: merged (idx [nq,k] i32 GLOBAL row ids, dist [nq,k] f64) over all shards."""
        q = np.ascontiguousarray(queries, dtype=np.float64).reshape(-1, dbs[0].dim)
        nq = q.shape[0]
        idx = np.full((nq, k), -1, np.int32)
        dist = np.full((nq, k), np.inf, np.float64)
        self.check(lib().sepb_group_netvlad_topk(self._h, self._handles(dbs), _ptr(q), nq, k, mode, _ptr(idx), _ptr(dist)))
        return idx, dist

    def netvlad_topk_dev(self, dbs, d_queries, nq, k, mode, d_out_idx, d_out_dist):
        """Device form: lists of device pointers (ints), one per local device; asynchronous on each context's stream."""
        n = len(self.ctxs)
        arr = lambda ps: (C.c_void_p * n)(*[C.c_void_p(int(p)) for p in ps])
        self.check(lib().sepb_group_netvlad_topk_dev(self._h, self._handles(dbs), arr(d_queries), nq, k, mode,
                                                     arr(d_out_idx), arr(d_out_dist)))

    def close(self):
        if self._h:
            for c in self.ctxs:
                for child in list(c._children):
                    child.close()
            lib().sepb_group_destroy(self._h)
            self._h = C.c_void_p()
            for c in self.ctxs:
                if not c._owned:
                    c._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

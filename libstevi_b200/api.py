"""Host-side mirror of the reference's ``StereoVision::Correlation`` functions over the C ABI.

Function names, argument order and meaning follow the reference templates (file:line cited per function);
template arguments become keyword arguments. Operands are numpy arrays (host memory: staged by the library) or
torch CUDA tensors (device memory: used in place, asynchronous on the context's stream). Results come back in the
same memory space as the inputs, shaped like the reference's and — with ``layout="ref"`` — laid out like them.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("STEVI_B200_LIB") or os.path.join(_HERE, "libstevi_b200.so")  # env override: kernel experiments

# enums (include/stevi_b200/c_api.h)
OK, E_SHAPE, E_DTYPE, E_UNSUPPORTED, E_CUDA, E_NCCL, E_ARG = range(7)
U8, U16, I32, U32, F32 = range(5)
HOST, DEVICE = 0, 1
COMPAT, CORRECTED = 0, 1
EXACT_SUMS = 256
CC, NCC, SSD, SAD, ZCC, ZNCC, ZSSD, ZSAD, MEDAD, ZMEDAD, HAMMING, CENSUS, KERMI = range(13)
COST, SCORE = 0, 1
LEFT_TO_RIGHT, RIGHT_TO_LEFT = 0, 1
TCV_SAME, TCV_REVERSED, TCV_BOTH = 0, 1, 2
EQUIANGULAR, PARABOLA, GAUSSIAN = 0, 1, 2
ISOTROPIC, ANISOTROPIC = 0, 1  # IsotropyHypothesis, cost_based_refinement.h:37-41
QUEEN, ROOK, BISHOP = 0, 1, 2  # Contiguity::bidimensionalContiguity, utils/contiguity.h:31-35
SCORE_FUNCS = (CC, NCC, ZCC, ZNCC)

_NP2DT = {np.dtype(np.uint8): U8, np.dtype(np.uint16): U16, np.dtype(np.int32): I32, np.dtype(np.uint32): U32,
          np.dtype(np.float32): F32}
_DT2NP = {v: k for k, v in _NP2DT.items()}


class SteviError(RuntimeError):
    def __init__(self, status, text):
        super().__init__(f"stevi status {status}: {text}")
        self.status = status


class StArray(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("dtype", ctypes.c_int32), ("ndim", ctypes.c_int32),
                ("shape", ctypes.c_int32 * 4), ("strides", ctypes.c_int64 * 4), ("mem", ctypes.c_int32)]


class StMargins(ctypes.Structure):
    _fields_ = [("left", ctypes.c_int32), ("top", ctypes.c_int32), ("right", ctypes.c_int32), ("bottom", ctypes.c_int32)]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(make -C libstevi_b200/csrc). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    lib.stevi_last_error.restype = ctypes.c_char_p
    lib.stevi_version.restype = ctypes.c_char_p
    lib.stevi_ctx_stream.restype = ctypes.c_void_p
    lib.stevi_ctx_timing_report.restype = ctypes.c_char_p
    lib.stevi_ctx_launch_count.restype = ctypes.c_int64
    lib.stevi_ctx_workspace_bytes.restype = ctypes.c_int64
    return lib


lib = _load()


def _is_torch(x):
    return hasattr(x, "data_ptr") and hasattr(x, "stride")


def _torch_dtype_code(t):
    import torch
    return {torch.uint8: U8, torch.uint16: U16, torch.int32: I32, torch.uint32: U32, torch.float32: F32}[t.dtype]


def describe(x):
    """stevi_array descriptor of a numpy array (host) or torch CUDA tensor (device). Keeps no reference."""
    a = StArray()
    if _is_torch(x):
        if not x.is_cuda:
            raise TypeError("torch operands must live on a CUDA device; pass numpy arrays for host memory")
        a.data = x.data_ptr()
        a.dtype = _torch_dtype_code(x)
        a.ndim = x.dim()
        for k in range(x.dim()):
            a.shape[k] = x.shape[k]
            a.strides[k] = x.stride(k)
        a.mem = DEVICE
    else:
        if x.dtype not in _NP2DT:
            raise TypeError(f"unsupported dtype {x.dtype}")
        a.data = x.ctypes.data
        a.dtype = _NP2DT[x.dtype]
        a.ndim = x.ndim
        for k in range(x.ndim):
            a.shape[k] = x.shape[k]
            if x.strides[k] % x.itemsize:
                raise ValueError("strides must be multiples of the item size")
            a.strides[k] = x.strides[k] // x.itemsize
        a.mem = HOST
    return a


class Context:
    """One per host thread / CUDA stream (stevi_ctx). ``stream``: a raw cudaStream_t value, or None for an own stream;
    ``torch_stream=True`` binds to torch's current stream on that device so torch events and tensors interoperate."""

    def __init__(self, device=0, stream=None, torch_stream=False):
        if torch_stream:
            import torch
            stream = torch.cuda.current_stream(device).cuda_stream
            if not stream:
                stream = 1  # torch's default stream is the legacy stream: cudaStreamLegacy == (cudaStream_t)0x1
        self._h = ctypes.c_void_p()
        rc = lib.stevi_ctx_create(int(device), ctypes.c_void_p(stream), ctypes.byref(self._h))
        if rc != OK:
            raise SteviError(rc, "stevi_ctx_create failed: no usable CUDA device (there is no CPU fallback)")
        self.device = device

    def close(self):
        if self._h:
            lib.stevi_ctx_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc):
        if rc != OK:
            raise SteviError(rc, lib.stevi_last_error(self._h).decode())

    def synchronize(self):
        self.check(lib.stevi_ctx_synchronize(self._h))

    def set_deferred_host(self, enable):
        """Calls with host operands enqueue their copies and return without waiting; results are valid after
        ``synchronize()`` (see stevi_ctx_set_deferred_host)."""
        self.check(lib.stevi_ctx_set_deferred_host(self._h, 1 if enable else 0))

    @property
    def stream(self):
        """The ctx's cudaStream_t as an integer."""
        return lib.stevi_ctx_stream(self._h) or 0

    def set_narrow_wire(self, enable):
        """Index / disparity maps returned to host memory cross PCIe as u8 / u16 and are widened by the host (default on)."""
        self.check(lib.stevi_ctx_set_narrow_wire(self._h, 1 if enable else 0))

    def set_row_origin(self, row):
        """Global frame row of row 0 of the images passed next (row-band runs; see stevi_ctx_set_row_origin)."""
        self.check(lib.stevi_ctx_set_row_origin(self._h, int(row)))

    def set_timing(self, enable):
        self.check(lib.stevi_ctx_set_timing(self._h, 1 if enable else 0))

    def timing_report(self):
        """[(kernel name, ms)] per launch since the previous report (synchronises the stream)."""
        txt = lib.stevi_ctx_timing_report(self._h).decode()
        return [(ln.split("\t")[0], float(ln.split("\t")[1])) for ln in txt.splitlines() if ln]

    @property
    def launches(self):
        return lib.stevi_ctx_launch_count(self._h)

    @property
    def workspace_bytes(self):
        return lib.stevi_ctx_workspace_bytes(self._h)

    @property
    def handle(self):
        return self._h


class HostPairStream:
    """Keeps ``depth`` stereo pairs in flight from page-locked host memory: slot k has its own Context (own stream, own
    workspace) with deferred host completion, so the host->device copy of one pair, the kernels of another and the
    device->host copy of a third overlap. The reference processes one pair per call on the host cores
    (benchmarkCostVolumeWorkflow.cpp:175-260 loops over pairs); this is the same loop with the calls left in flight.

        stream = HostPairStream(depth=3)
        for k, (left, right) in enumerate(pairs):        # pinned numpy arrays
            ctx = stream.acquire()                        # waits for the pair that used this slot `depth` calls ago
            censusSgmPipeline(left, right, ..., out=results[k % 3], ctx=ctx)
        stream.drain()
    """

    def __init__(self, depth=3, device=0, narrow_wire=True):
        if depth < 1:
            raise ValueError("depth must be positive")
        self.slots = [Context(device) for _ in range(depth)]
        for c in self.slots:
            c.set_deferred_host(True)
            c.set_narrow_wire(narrow_wire)
        self._next = 0

    @property
    def slot(self):
        """Index of the slot the next ``acquire`` hands out (callers key their result buffers on it)."""
        return self._next

    def acquire(self):
        ctx = self.slots[self._next]
        self._next = (self._next + 1) % len(self.slots)
        ctx.synchronize()
        return ctx

    def drain(self):
        for c in self.slots:
            c.synchronize()

    @property
    def launches(self):
        return sum(c.launches for c in self.slots)

    def close(self):
        for c in self.slots:
            c.close()


_default_ctx = {}


def default_context(device=0):
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]


_default_torch_ctx = {}


def _ctx_for(ctx, x):
    """The context a call without ``ctx=`` runs on. Host operands: an own-stream context (the call returns with its results in
    place). Torch CUDA tensors: a context bound to torch's CURRENT stream on the tensor's device, re-bound on every call, so
    the kernels are ordered after whatever produced the operands and before whatever reads the results, and the caching
    allocator's stream bookkeeping holds. (A ``Context()`` with its own stream is for callers who order things themselves.)"""
    if ctx is not None:
        return ctx
    if not _is_torch(x):
        return default_context(0)
    import torch
    dev = x.device.index
    stream = torch.cuda.current_stream(dev).cuda_stream or 1  # legacy default stream == (cudaStream_t)0x1
    c = _default_torch_ctx.get(dev)
    if c is None:
        c = _default_torch_ctx[dev] = Context(dev, stream=stream)
    else:
        c.check(lib.stevi_ctx_set_stream(c.handle, ctypes.c_void_p(stream)))
    return c


def _empty(like, shape, np_dtype, order=None):
    """Uninitialised array in the memory space of ``like``; ``order`` = permutation giving the memory order
    (slowest axis first), default C order."""
    shape = tuple(int(s) for s in shape)
    order = tuple(range(len(shape))) if order is None else tuple(order)
    mem_shape = tuple(shape[k] for k in order)
    inv = tuple(order.index(k) for k in range(len(shape)))
    if _is_torch(like):
        import torch
        tdt = {np.dtype(np.uint8): torch.uint8, np.dtype(np.uint16): torch.uint16, np.dtype(np.int32): torch.int32,
               np.dtype(np.uint32): torch.uint32, np.dtype(np.float32): torch.float32}[np.dtype(np_dtype)]
        return torch.empty(mem_shape, dtype=tdt, device=like.device).permute(inv)
    return np.empty(mem_shape, dtype=np_dtype).transpose(inv)


# memory orders of (H,W,D) volumes
ORDER_C = (0, 1, 2)      # [H][W][D]
ORDER_REF_CV = (0, 2, 1)  # [H][D][W]: aggregateCost's strides {W*D, 1, W} (cross_correlations.h:220)


def _volume_order(layout):
    if layout in ("ref", "hdw"):
        return ORDER_REF_CV
    if layout in ("c", "hwd"):
        return ORDER_C
    if layout in ("f", "dwh"):
        return (2, 1, 0)  # first index fastest
    raise ValueError(layout)


def strategy_of(matchfunc):
    return SCORE if matchfunc in SCORE_FUNCS else COST


def censusTransform2D(img, h_radius, v_radius, semantics=COMPAT, ctx=None):
    """census.h:117-131. (H,W) u8|f32 -> (H,W,nwords) u32."""
    ctx = _ctx_for(ctx, img)
    F = (2 * h_radius + 1) * (2 * v_radius + 1)
    out = _empty(img, img.shape + ((F - 1) // 32 + 1,), np.uint32)
    a, o = describe(img), describe(out)
    ctx.check(lib.stevi_census_transform(ctx.handle, ctypes.byref(a), h_radius, v_radius, semantics, ctypes.byref(o)))
    return out


def unfoldBasedCostVolume(matchFunc, img_l, img_r, h_radius, v_radius, disp_width, dDir=RIGHT_TO_LEFT, TCV=None,
                          semantics=COMPAT, layout="ref", exact_sums=False, ctx=None):
    """cross_correlations.h:740-765. Returns None where the reference returns an empty array.
    ``exact_sums``: float matching functions evaluate the reference's O(F) per-cell sums instead of the box sums."""
    ctx = _ctx_for(ctx, img_l)
    if exact_sums:
        semantics |= EXACT_SUMS
    if TCV is None:
        TCV = np.uint16 if matchFunc in (CENSUS, HAMMING) else np.float32
    ws = img_r.shape[1] if dDir == RIGHT_TO_LEFT else img_l.shape[1]
    out = _empty(img_l, (img_l.shape[0], ws, disp_width), TCV, _volume_order(layout))
    l, r, o = describe(img_l), describe(img_r), describe(out)
    rc = lib.stevi_cost_volume(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), h_radius, v_radius, disp_width,
                               dDir, semantics, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def hammingCostVolume(feat_l, feat_r, disp_width, dDir=RIGHT_TO_LEFT, TCV=np.uint16, semantics=COMPAT, layout="ref", ctx=None):
    """aggregateCost<HAMMING> on (H,W,nw) u32 features, cross_correlations.h:194-249."""
    ctx = _ctx_for(ctx, feat_l)
    ws = feat_r.shape[1] if dDir == RIGHT_TO_LEFT else feat_l.shape[1]
    out = _empty(feat_l, (feat_l.shape[0], ws, disp_width), TCV, _volume_order(layout))
    l, r, o = describe(feat_l), describe(feat_r), describe(out)
    rc = lib.stevi_hamming_cost_volume(ctx.handle, ctypes.byref(l), ctypes.byref(r), disp_width, dDir, semantics, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


NOALLOC_SCRATCH_BYTES = 16 << 20   # STEVI_NOALLOC_SCRATCH_BYTES (c_api.h)


def censusCostVolumeNoAlloc(img_l, img_r, h_radius, v_radius, disp_width, out, dDir=RIGHT_TO_LEFT, semantics=COMPAT, ctx=None):
    """stevi_census_cost_volume_noalloc: unfoldBasedCostVolume<CENSUS> into the caller's ``out`` with the census words in the
    library's static device scratch array (no counterpart in the reference; for callers that cannot call cudaMalloc). Device
    tensors only; nothing is allocated here either: ``out`` (H, Wsource, D) u16|i32|f32 comes from the caller and is returned.
    Pairs whose census words exceed NOALLOC_SCRATCH_BYTES raise SteviError; calls that overlap on the GPU race on the array."""
    ctx = _ctx_for(ctx, img_l)
    l, r, o = describe(img_l), describe(img_r), describe(out)
    ctx.check(lib.stevi_census_cost_volume_noalloc(ctx.handle, ctypes.byref(l), ctypes.byref(r), h_radius, v_radius, disp_width,
                                                   dDir, semantics, ctypes.byref(o)))
    return out


def getFeatureVolumeForMatchFuncFromFeatures(matchFunc, feature_vol, ctx=None):
    """getFeatureVolumeForMatchFunc<matchFunc>(feature_vol) on a ready-made (H,W,F) f32 volume, cross_correlations.h:645-722."""
    ctx = _ctx_for(ctx, feature_vol)
    out = _empty(feature_vol, feature_vol.shape, np.float32)
    a, o = describe(feature_vol), describe(out)
    ctx.check(lib.stevi_features_for_matchfunc(ctx.handle, matchFunc, ctypes.byref(a), ctypes.byref(o)))
    return out


def aggregateCost(matchFunc, feat_l, feat_r, disp_width=None, searchRange=None, dDir=RIGHT_TO_LEFT, layout="ref", ctx=None):
    """aggregateCost on processed float feature volumes, cross_correlations.h:194-249; ``searchRange=(lower, upper)`` is the
    searchOffset<1> overload (:251-308). Returns None where the reference returns an empty array."""
    ctx = _ctx_for(ctx, feat_l)
    lower, upper = (0, disp_width - 1) if searchRange is None else searchRange
    ws = feat_r.shape[1] if dDir == RIGHT_TO_LEFT else feat_l.shape[1]
    if upper < lower:
        return None
    out = _empty(feat_l, (feat_l.shape[0], ws, upper - lower + 1), np.float32, _volume_order(layout))
    l, r, o = describe(feat_l), describe(feat_r), describe(out)
    rc = lib.stevi_aggregate_cost(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), lower, upper, dDir, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def featureVolume2CostVolume(matchFunc, feature_vol_l, feature_vol_r, disp_width, dDir=RIGHT_TO_LEFT, TCV=None, semantics=COMPAT,
                             layout="ref", ctx=None):
    """cross_correlations.h:724-738: getFeatureVolumeForMatchFunc of both volumes, then aggregateCost."""
    ctx = _ctx_for(ctx, feature_vol_l)
    if TCV is None:
        TCV = np.uint16 if matchFunc in (CENSUS, HAMMING) else np.float32
    ws = feature_vol_r.shape[1] if dDir == RIGHT_TO_LEFT else feature_vol_l.shape[1]
    out = _empty(feature_vol_l, (feature_vol_l.shape[0], ws, disp_width), TCV, _volume_order(layout))
    l, r, o = describe(feature_vol_l), describe(feature_vol_r), describe(out)
    rc = lib.stevi_feature_volume_to_cost_volume(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), disp_width, dDir, semantics,
                                                 ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def unfoldBasedCostVolumeCompressed(matchFunc, img_l, img_r, compressor_mask, disp_width, dDir=RIGHT_TO_LEFT, TCV=None,
                                    semantics=COMPAT, layout="ref", ctx=None):
    """unfoldBasedCostVolume(img_l, img_r, UnFoldCompressor(mask), disp_width), cross_correlations.h:767-791. Images (H,W) or (H,W,C)."""
    ctx = _ctx_for(ctx, img_l)
    if TCV is None:
        TCV = np.uint16 if matchFunc in (CENSUS, HAMMING) else np.float32
    mask = np.ascontiguousarray(compressor_mask, np.int32)
    ws = img_r.shape[1] if dDir == RIGHT_TO_LEFT else img_l.shape[1]
    out = _empty(img_l, (img_l.shape[0], ws, disp_width), TCV, _volume_order(layout))
    m, l, r, o = describe(mask), describe(img_l), describe(img_r), describe(out)
    rc = lib.stevi_cost_volume_compressed(ctx.handle, matchFunc, ctypes.byref(m), ctypes.byref(l), ctypes.byref(r), disp_width, dDir,
                                          semantics, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def unfoldBasedCostVolumeMultiChannel(matchFunc, img_l, img_r, h_radius, v_radius, disp_width, dDir=RIGHT_TO_LEFT, TCV=None,
                                      semantics=COMPAT, layout="ref", ctx=None):
    """unfoldBasedCostVolume<matchFunc, T, T, 3> on (H,W,C) images, cross_correlations.h:740-765 with unfold.h:293-344."""
    ctx = _ctx_for(ctx, img_l)
    if TCV is None:
        TCV = np.uint16 if matchFunc in (CENSUS, HAMMING) else np.float32
    ws = img_r.shape[1] if dDir == RIGHT_TO_LEFT else img_l.shape[1]
    out = _empty(img_l, (img_l.shape[0], ws, disp_width), TCV, _volume_order(layout))
    l, r, o = describe(img_l), describe(img_r), describe(out)
    rc = lib.stevi_cost_volume_multichannel(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), h_radius, v_radius, disp_width,
                                            dDir, semantics, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def sgmCostVolume(cv, P1, P2, margins=(0, 0, 0, 0), Pout=100.0, nDirections=8, strategy=COST, semantics=COMPAT,
                  layout="same", ctx=None):
    """sgm.h:360-404. margins = (left, top, right, bottom) like StereoVision::Margins' 4-argument constructor."""
    ctx = _ctx_for(ctx, cv)
    if layout == "same":
        if _is_torch(cv):
            order = tuple(sorted(range(3), key=lambda k: -cv.stride(k)))
        else:
            order = tuple(sorted(range(3), key=lambda k: -cv.strides[k]))
    else:
        order = _volume_order(layout)
    out = _empty(cv, cv.shape, np.float32, order)
    m = StMargins(*[int(v) for v in margins])
    a, o = describe(cv), describe(out)
    f32 = ctypes.c_float
    ctx.check(lib.stevi_sgm(ctx.handle, ctypes.byref(a), nDirections, strategy, f32(P1), f32(P2), f32(Pout),
                            ctypes.byref(m), semantics, ctypes.byref(o)))
    return out


def extractSelectedIndex(vol, strategy=COST, ctx=None):
    """correlation_base.h:427-464."""
    ctx = _ctx_for(ctx, vol)
    out = _empty(vol, vol.shape[:2], np.int32)
    a, o = describe(vol), describe(out)
    ctx.check(lib.stevi_wta(ctx.handle, ctypes.byref(a), strategy, ctypes.byref(o)))
    return out


def selectedIndexToDisp(idx, offset=0, dDir=RIGHT_TO_LEFT, ctx=None):
    """correlation_base.h:511-532."""
    ctx = _ctx_for(ctx, idx)
    out = _empty(idx, idx.shape, np.int32)
    a, o = describe(idx), describe(out)
    ctx.check(lib.stevi_index_to_disp(ctx.handle, ctypes.byref(a), dDir, offset, ctypes.byref(o)))
    return out


def _search4(search):
    lo0, up0, lo1, up1 = [int(v) for v in search]
    return (ctypes.c_int32 * 4)(lo0, up0, lo1, up1), up0 - lo0 + 1, up1 - lo1 + 1


def unfoldBased2dDisparityCostVolume(matchFunc, img_l, img_r, h_radius, v_radius, searchWindows, dDir=RIGHT_TO_LEFT, layout="ref",
                                     exact_sums=False, ctx=None):
    """cross_correlations.h:794-822. ``searchWindows`` = (lower0, upper0, lower1, upper1), the arguments of searchOffset<2>
    (axis 0 = rows, axis 1 = columns). Returns (H, W, dH, dW) float32; ``layout`` "ref" = the reference's strides
    {W*dW*dH, 1, W*dW, W}, "c" = C order. ``exact_sums``: every cell summed over its features in the reference's channel order
    (2e-6) instead of the O(1) box sums (1e-5). Raises SteviError(E_SHAPE) where the reference returns an empty array."""
    ctx = _ctx_for(ctx, img_l)
    so, dH, dW = _search4(searchWindows)
    H, W = img_l.shape
    out = _empty(img_l, (H, W, max(dH, 0), max(dW, 0)), np.float32, (0, 2, 3, 1) if layout == "ref" else None)
    l, r, o = describe(img_l), describe(img_r), describe(out)
    ctx.check(lib.stevi_cost_volume_2d_ex(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), h_radius, v_radius, so, dDir,
                                          EXACT_SUMS if exact_sums else COMPAT, ctypes.byref(o)))
    return out


def extractSelected2dIndex(vol, strategy=COST, ctx=None):
    """correlation_base.h:466-508: (H, W, D1, D2) -> (H, W, 2) int32."""
    ctx = _ctx_for(ctx, vol)
    out = _empty(vol, tuple(vol.shape[:2]) + (2,), np.int32)
    a, o = describe(vol), describe(out)
    ctx.check(lib.stevi_wta_2d(ctx.handle, ctypes.byref(a), strategy, ctypes.byref(o)))
    return out


def selected2dIndexToDisp(idx, searchWindows, ctx=None):
    """correlation_base.h:534-555."""
    ctx = _ctx_for(ctx, idx)
    so, _, _ = _search4(searchWindows)
    out = _empty(idx, idx.shape, np.int32)
    a, o = describe(idx), describe(out)
    ctx.check(lib.stevi_index2d_to_disp(ctx.handle, ctypes.byref(a), so, ctypes.byref(o)))
    return out


def truncatedBidirectionaCostVolume(cv, idx, cost_vol_radius0=255, cost_vol_radius1=255, ctx=None):
    """correlation_base.h:677-725 (sic). (H, W, D1, D2) f32 + (H, W, 2) i32 -> (H, W, 2*r0+1, 2*r1+1) f32, NaN outside the
    search window. The radii are the reference's uint8_t arguments: its default uint8_t(-1) is 255; 0 means 1 for axis 0
    and "same as axis 0" for axis 1."""
    ctx = _ctx_for(ctx, cv)
    r0 = max(int(cost_vol_radius0), 1)
    r1 = int(cost_vol_radius1) if int(cost_vol_radius1) >= 1 else r0
    out = _empty(cv, tuple(cv.shape[:2]) + (2 * r0 + 1, 2 * r1 + 1), np.float32)
    a, i, o = describe(cv), describe(idx), describe(out)
    ctx.check(lib.stevi_truncated_cost_volume_2d(ctx.handle, ctypes.byref(a), ctypes.byref(i), cost_vol_radius0, cost_vol_radius1, ctypes.byref(o)))
    return out


def refineDisp2dCostInterpolation(tcv, rawDisparity, kernel=PARABOLA, isotropHypothesis=ISOTROPIC, ctx=None):
    """cost_based_refinement.h:165-369. (H, W, T0, T1) f32 + (H, W, 2) i32 -> (H, W, 2) f32. Returns None where the reference
    returns an empty array (a side of the neighbourhood even or < 3)."""
    ctx = _ctx_for(ctx, tcv)
    out = _empty(tcv, tuple(tcv.shape[:2]) + (2,), np.float32)
    t, i, o = describe(tcv), describe(rawDisparity), describe(out)
    rc = lib.stevi_refine_disp_2d_cost_interp(ctx.handle, ctypes.byref(t), ctypes.byref(i), kernel, isotropHypothesis, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def refineDisp2dCostPatchInterpolation(tcv, rawDisparity, kernel=PARABOLA, ctx=None):
    """cost_based_refinement.h:371-433: stationary point of the least-squares quadric through the 3x3 centre cells.
    Returns None where the reference returns an empty array; Equiangular is rejected (the reference static_asserts)."""
    ctx = _ctx_for(ctx, tcv)
    out = _empty(tcv, tuple(tcv.shape[:2]) + (2,), np.float32)
    t, i, o = describe(tcv), describe(rawDisparity), describe(out)
    rc = lib.stevi_refine_disp_2d_cost_patch_interp(ctx.handle, ctypes.byref(t), ctypes.byref(i), kernel, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def computeAggregationWindows(guide, h_radius, v_radius, spatialWeightFunc, matchWeightFunc):
    """guided_cost_filtering.h:26-100: weights(i,j,di+v,dj+h) = spatialWeightFunc(di, dj) * matchWeightFunc(guide(i,j,:), guide(i+di,j+dj,:)),
    0 for window cells outside the image. The two functors are the caller's code and run on the host: ``spatialWeightFunc(di, dj)``
    returns a number, ``matchWeightFunc(ref, target)`` takes two (..., C) numpy arrays and returns the weights of all pixels at once.
    The result feeds variableCostVolumeAggregation, which is where the volume-sized work is."""
    g = np.asarray(guide)
    if g.ndim == 2:
        g = g[:, :, None]
    H, W = g.shape[:2]
    out = np.zeros((H, W, 2 * v_radius + 1, 2 * h_radius + 1), np.float32)
    for di in range(-v_radius, v_radius + 1):
        for dj in range(-h_radius, h_radius + 1):
            i0, i1, j0, j1 = max(0, -di), min(H, H - di), max(0, -dj), min(W, W - dj)
            if i0 >= i1 or j0 >= j1:
                continue
            w = np.float32(spatialWeightFunc(di, dj)) * np.asarray(matchWeightFunc(g[i0:i1, j0:j1], g[i0 + di:i1 + di, j0 + dj:j1 + dj]), np.float32)
            out[i0:i1, j0:j1, di + v_radius, dj + h_radius] = w
    return out


def variableCostVolumeAggregation(costVolume, aggregationWindows, layout="c", out=None, ctx=None):
    """guided_cost_filtering.h:102-176. (H, W, D) f32 volume + (H, W, 2v+1, 2h+1) f32 per-pixel weights -> (H, W, D) f32 weighted
    means over each pixel's (image-clipped) window. ``layout`` "ref" gives the result the strides {W*D, 1, W} of the reference's
    cost volumes, "c" = C order (what the reference's ``Array(shape)`` gives). Returns None where the reference returns an empty
    array (windows of another size, even or empty window sides)."""
    ctx = _ctx_for(ctx, costVolume)
    if out is None:
        out = _empty(costVolume, tuple(costVolume.shape), np.float32, (0, 2, 1) if layout == "ref" else None)
    a, w, o = describe(costVolume), describe(aggregationWindows), describe(out)
    rc = lib.stevi_variable_cost_volume_aggregation(ctx.handle, ctypes.byref(a), ctypes.byref(w), ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def censusSgmPipeline(img_l, img_r, h_radius, v_radius, disp_width, P1, P2, Pout=100.0, nDirections=8,
                      margins=(0, 0, 0, 0), dDir=RIGHT_TO_LEFT, semantics=COMPAT, return_volumes=False, layout="ref",
                      out=None, ctx=None):
    """Census cost volume -> SGM -> WTA -> disparity in one call (benchmarkCrossCorrelationAlgorithms.cpp:289-293 with
    CENSUS costs). Returns disp, or (cv, sgm, disp) with ``return_volumes``. ``out`` = preallocated (cv, sgm, disp) or disp."""
    ctx = _ctx_for(ctx, img_l)
    H = img_l.shape[0]
    ws = img_r.shape[1] if dDir == RIGHT_TO_LEFT else img_l.shape[1]
    cv = sg = None
    if out is not None:
        if isinstance(out, tuple):
            cv, sg, disp = out
        else:
            disp = out
    else:
        disp = _empty(img_l, (H, ws), np.int32)
        if return_volumes:
            cv = _empty(img_l, (H, ws, disp_width), np.uint16, _volume_order(layout))
            sg = _empty(img_l, (H, ws, disp_width), np.float32, _volume_order(layout))
    m = StMargins(*[int(v) for v in margins])
    l, r, d = describe(img_l), describe(img_r), describe(disp)
    dcv = describe(cv) if cv is not None else None
    dsg = describe(sg) if sg is not None else None
    f32 = ctypes.c_float
    ctx.check(lib.stevi_census_sgm_pipeline(
        ctx.handle, ctypes.byref(l), ctypes.byref(r), h_radius, v_radius, disp_width, dDir, nDirections,
        f32(P1), f32(P2), f32(Pout), ctypes.byref(m), semantics,
        ctypes.byref(dcv) if dcv is not None else None, ctypes.byref(dsg) if dsg is not None else None, ctypes.byref(d)))
    if cv is not None or sg is not None:
        return cv, sg, disp
    return disp


class StPair(ctypes.Structure):
    _fields_ = [("img_l", ctypes.POINTER(StArray)), ("img_r", ctypes.POINTER(StArray)), ("disp", ctypes.POINTER(StArray))]


def batchCensusSgm(ctxs, pairs, h_radius, v_radius, disp_width, P1, P2, Pout=100.0, nDirections=8, margins=(0, 0, 0, 0),
                   dDir=RIGHT_TO_LEFT, semantics=COMPAT):
    """stevi_batch_census_sgm: ``pairs`` = [(img_l, img_r, disp_out)], ``ctxs`` = contexts (own streams) on one or several devices;
    the library keeps one pair in flight per context and drives every device from its own host thread. Results land in disp_out."""
    n = len(pairs)
    descs = [(describe(l), describe(r), describe(d)) for l, r, d in pairs]
    arr = (StPair * max(n, 1))()
    for k, (l, r, d) in enumerate(descs):
        arr[k].img_l, arr[k].img_r, arr[k].disp = ctypes.pointer(l), ctypes.pointer(r), ctypes.pointer(d)
    hs = (ctypes.c_void_p * len(ctxs))(*[c.handle for c in ctxs])
    m = StMargins(*[int(v) for v in margins])
    f32 = ctypes.c_float
    rc = lib.stevi_batch_census_sgm(hs, len(ctxs), arr, n, h_radius, v_radius, disp_width, dDir, nDirections, f32(P1), f32(P2), f32(Pout),
                                    ctypes.byref(m), semantics)
    if rc != OK:
        texts = [lib.stevi_last_error(c.handle).decode() for c in ctxs]
        raise SteviError(rc, "; ".join(t for t in texts if t))


def selectedCost(cv, idx, ctx=None):
    """correlation_base.h:557-577: out(i,j) = cv(i,j,idx(i,j))."""
    ctx = _ctx_for(ctx, cv)
    np_dt = _DT2NP[_torch_dtype_code(cv)] if _is_torch(cv) else cv.dtype
    out = _empty(cv, cv.shape[:2], np_dt)
    a, i, o = describe(cv), describe(idx), describe(out)
    ctx.check(lib.stevi_selected_cost(ctx.handle, ctypes.byref(a), ctypes.byref(i), ctypes.byref(o)))
    return out


def _tcv_depth(radius, sdir):
    return 4 * radius + 1 if sdir == TCV_BOTH else 2 * radius + 1


def truncatedCostVolume(cv, idx, h_radius, v_radius, cost_vol_radius, dDir=RIGHT_TO_LEFT, sdir=TCV_SAME, ctx=None):
    """correlation_base.h:579-674. (H,W,D) f32 + (H,W) i32 -> (H,W,2r+1 | 4r+1) f32 with NaN outside."""
    ctx = _ctx_for(ctx, cv)
    out = _empty(cv, tuple(cv.shape[:2]) + (_tcv_depth(cost_vol_radius, sdir),), np.float32)
    a, i, o = describe(cv), describe(idx), describe(out)
    ctx.check(lib.stevi_truncated_cost_volume(ctx.handle, ctypes.byref(a), ctypes.byref(i), h_radius, v_radius, cost_vol_radius,
                                              dDir, sdir, ctypes.byref(o)))
    return out


def extractInBoundDomain(idx, width, h_radius, v_radius, cost_vol_radius, dDir=RIGHT_TO_LEFT, sdir=TCV_SAME, ctx=None):
    """correlation_base.h:727-822. (H,W) i32 -> (H,W,2r+1 | 4r+1) f32 mask of 1/0."""
    ctx = _ctx_for(ctx, idx)
    out = _empty(idx, tuple(idx.shape) + (_tcv_depth(cost_vol_radius, sdir),), np.float32)
    i, o = describe(idx), describe(out)
    ctx.check(lib.stevi_in_bound_domain(ctx.handle, ctypes.byref(i), width, h_radius, v_radius, cost_vol_radius, dDir, sdir,
                                        ctypes.byref(o)))
    return out


def refineDispCostInterpolation(tcv, idx, kernel=PARABOLA, ctx=None):
    """cost_based_refinement.h:128-163. Returns None where the reference returns an empty array (even / too small depth)."""
    ctx = _ctx_for(ctx, tcv)
    out = _empty(tcv, tcv.shape[:2], np.float32)
    t, i, o = describe(tcv), describe(idx), describe(out)
    rc = lib.stevi_refine_disp_cost_interp(ctx.handle, ctypes.byref(t), ctypes.byref(i), kernel, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def _source_of(feat_l, feat_r, dDir):
    return feat_r if dDir == RIGHT_TO_LEFT else feat_l


def refineBarycentricDisp(matchFunc, feature_vol_l, feature_vol_r, selectedIndex, dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:184-355 (NCC, ZNCC, SSD, ZSSD, SAD, ZSAD). (H,W,F) f32 feature volumes, (H,W) i32 winners."""
    ctx = _ctx_for(ctx, feature_vol_l)
    out = _empty(feature_vol_l, _source_of(feature_vol_l, feature_vol_r, dDir).shape[:2], np.float32)
    l, r, i, o = describe(feature_vol_l), describe(feature_vol_r), describe(selectedIndex), describe(out)
    ctx.check(lib.stevi_refine_barycentric_disp(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), ctypes.byref(i), dDir, ctypes.byref(o)))
    return out


def refineBarycentricSymmetricDisp(matchFunc, feature_vol_l, feature_vol_r, selectedIndex, disp_width, refineRadius=1,
                                   dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:31-105 (NCC, ZNCC, SSD, ZSSD; refineRadius 1..3)."""
    ctx = _ctx_for(ctx, feature_vol_l)
    out = _empty(feature_vol_l, _source_of(feature_vol_l, feature_vol_r, dDir).shape[:2], np.float32)
    l, r, i, o = describe(feature_vol_l), describe(feature_vol_r), describe(selectedIndex), describe(out)
    ctx.check(lib.stevi_refine_barycentric_symmetric_disp(ctx.handle, matchFunc, refineRadius, ctypes.byref(l), ctypes.byref(r),
                                                          ctypes.byref(i), disp_width, dDir, ctypes.byref(o)))
    return out


def refineBarycentric2dDisp(matchFunc, feature_vol_l, feature_vol_r, selectedIndices, searchWindows, contiguity=QUEEN,
                            dDir=RIGHT_TO_LEFT, symmetric=False, ctx=None):
    """image_based_refinement.h:675-880; ``symmetric=True``: refineBarycentricSymmetric2dDisp, :1068-1241."""
    ctx = _ctx_for(ctx, feature_vol_l)
    out = _empty(feature_vol_l, tuple(feature_vol_l.shape[:2]) + (2,), np.float32)
    l, r, i, o = describe(feature_vol_l), describe(feature_vol_r), describe(selectedIndices), describe(out)
    ctx.check(lib.stevi_refine_barycentric_2d_disp(ctx.handle, matchFunc, contiguity, int(bool(symmetric)), ctypes.byref(l), ctypes.byref(r),
                                                   ctypes.byref(i), _search4(searchWindows)[0], dDir, ctypes.byref(o)))
    return out


def refineBarycentricSymmetric2dDisp(matchFunc, feature_vol_l, feature_vol_r, selectedIndices, searchWindows, contiguity=QUEEN,
                                     dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:1068-1241."""
    return refineBarycentric2dDisp(matchFunc, feature_vol_l, feature_vol_r, selectedIndices, searchWindows, contiguity, dDir, True, ctx)


def refineCostSymmetricDisp(matchFunc, feature_vol_l, feature_vol_r, selectedIndex, cost_volume, dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:1591-1690 (CC .. ZSAD)."""
    ctx = _ctx_for(ctx, feature_vol_l)
    out = _empty(feature_vol_l, _source_of(feature_vol_l, feature_vol_r, dDir).shape[:2], np.float32)
    l, r, i, c, o = (describe(feature_vol_l), describe(feature_vol_r), describe(selectedIndex), describe(cost_volume), describe(out))
    ctx.check(lib.stevi_refine_cost_symmetric_disp(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), ctypes.byref(i), ctypes.byref(c),
                                                   dDir, ctypes.byref(o)))
    return out


ZERO_MEAN_FUNCS = (ZCC, ZNCC, ZSSD, ZSAD)
NORMALIZED_FUNCS = (NCC, ZNCC)


def _refinement_volumes(matchFunc, feature_vol_l, feature_vol_r, preNormalize, ctx):
    """The volumes refined*FeatureVol (image_based_refinement.h:1692-1765, :1854-1927) matches on and refines on: the matching
    volumes are the function's processed features; the refinement runs on the zero-mean (not normalised) ones unless
    ``preNormalize``. Zero mean only = the processed features of the function's un-normalised sibling."""
    match = [getFeatureVolumeForMatchFuncFromFeatures(matchFunc, f, ctx=ctx) for f in (feature_vol_l, feature_vol_r)]
    if matchFunc not in NORMALIZED_FUNCS or preNormalize:
        return match, match
    if matchFunc in ZERO_MEAN_FUNCS:
        return match, [getFeatureVolumeForMatchFuncFromFeatures(ZCC, f, ctx=ctx) for f in (feature_vol_l, feature_vol_r)]
    return match, [feature_vol_l, feature_vol_r]


def refinedBarycentricDispFeatureVol(matchFunc, feature_vol_l, feature_vol_r, searchRange, preNormalize=False, dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:1854-1927: process the features, aggregateCost, extractSelectedIndex, refineBarycentricDisp."""
    ctx = _ctx_for(ctx, feature_vol_l)
    match, fine = _refinement_volumes(matchFunc, feature_vol_l, feature_vol_r, preNormalize, ctx)
    cv = aggregateCost(matchFunc, match[0], match[1], searchRange, dDir=dDir, ctx=ctx)
    if cv is None:
        return None
    idx = extractSelectedIndex(cv, strategy_of(matchFunc), ctx=ctx)
    return refineBarycentricDisp(matchFunc, fine[0], fine[1], idx, dDir, ctx=ctx)


def refinedBarycentricSymmetricDispFeatureVol(matchFunc, feature_vol_l, feature_vol_r, searchRange, preNormalize=False, refineRadius=1,
                                              dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:1692-1765."""
    ctx = _ctx_for(ctx, feature_vol_l)
    match, fine = _refinement_volumes(matchFunc, feature_vol_l, feature_vol_r, preNormalize, ctx)
    cv = aggregateCost(matchFunc, match[0], match[1], searchRange, dDir=dDir, ctx=ctx)
    if cv is None:
        return None
    idx = extractSelectedIndex(cv, strategy_of(matchFunc), ctx=ctx)
    return refineBarycentricSymmetricDisp(matchFunc, fine[0], fine[1], idx, searchRange, refineRadius, dDir, ctx=ctx)


def refinedCostSymmetricDispFeatureVol(matchFunc, feature_vol_l, feature_vol_r, searchRange, preNormalize=False, dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:2351-2424."""
    ctx = _ctx_for(ctx, feature_vol_l)
    match, fine = _refinement_volumes(matchFunc, feature_vol_l, feature_vol_r, preNormalize, ctx)
    cv = aggregateCost(matchFunc, match[0], match[1], searchRange, dDir=dDir, ctx=ctx)
    if cv is None:
        return None
    idx = extractSelectedIndex(cv, strategy_of(matchFunc), ctx=ctx)
    return refineCostSymmetricDisp(matchFunc, fine[0], fine[1], idx, cv, dDir, ctx=ctx)


def refinedBarycentricDisp(matchFunc, img_l, img_r, h_radius, v_radius, searchRange, preNormalize=False, dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:2502-2544: unfold both images, then refinedBarycentricDispFeatureVol. None where the reference
    returns an empty array (heights differ)."""
    if img_l.shape[0] != img_r.shape[0]:
        return None
    ctx = _ctx_for(ctx, img_l)
    fl, fr = (unfold(h_radius, v_radius, im, T_O=np.float32, ctx=ctx) for im in (img_l, img_r))
    return refinedBarycentricDispFeatureVol(matchFunc, fl, fr, searchRange, preNormalize, dDir, ctx=ctx)


def refinedBarycentricSymmetricDisp(matchFunc, img_l, img_r, h_radius, v_radius, searchRange, preNormalize=False, refineRadius=1,
                                    dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:2426-2470."""
    if img_l.shape[0] != img_r.shape[0]:
        return None
    ctx = _ctx_for(ctx, img_l)
    fl, fr = (unfold(h_radius, v_radius, im, T_O=np.float32, ctx=ctx) for im in (img_l, img_r))
    return refinedBarycentricSymmetricDispFeatureVol(matchFunc, fl, fr, searchRange, preNormalize, refineRadius, dDir, ctx=ctx)


def refinedCostSymmetricDisp(matchFunc, img_l, img_r, h_radius, v_radius, searchRange, preNormalize=False, dDir=RIGHT_TO_LEFT, ctx=None):
    """image_based_refinement.h:2573-2598."""
    if img_l.shape[0] != img_r.shape[0]:
        return None
    ctx = _ctx_for(ctx, img_l)
    fl, fr = (unfold(h_radius, v_radius, im, T_O=np.float32, ctx=ctx) for im in (img_l, img_r))
    return refinedCostSymmetricDispFeatureVol(matchFunc, fl, fr, searchRange, preNormalize, dDir, ctx=ctx)


def _search_arg(searchOffset):
    v = [int(x) for x in searchOffset]
    if len(v) not in (2, 4):
        raise ValueError("searchOffset: (lower, upper) for a 1-D search, (lower0, upper0, lower1, upper1) for a 2-D one")
    return (ctypes.c_int32 * len(v))(*v), len(v) // 2


def patchMatch(matchFunc, feature_vol_s, feature_vol_t, searchOffset, nIter=5, nRandomSearch=4, initializer=None, seed=0, ctx=None):
    """patchMatch<matchFunc, searchSpaceDim>, correlation/patchmatch.h:498-556: (H,W,F) f32 RAW feature volumes (source, target),
    ``searchOffset`` = (lower, upper) or (lower0, upper0, lower1, upper1) -> disparities (H,W,searchSpaceDim) int32. ``initializer``:
    an (H,W,searchSpaceDim) int32 map to start from (the reference's initializer functor's result); None = uniform random in the
    range. ``seed`` keys the counter-based generator of the random draws. None where the reference returns an empty array."""
    ctx = _ctx_for(ctx, feature_vol_s)
    so, dims = _search_arg(searchOffset)
    out = _empty(feature_vol_s, tuple(feature_vol_s.shape[:2]) + (dims,), np.int32)
    s, t, o = describe(feature_vol_s), describe(feature_vol_t), describe(out)
    i = describe(initializer) if initializer is not None else None
    rc = lib.stevi_patch_match(ctx.handle, matchFunc, ctypes.byref(s), ctypes.byref(t), dims, so, nIter, nRandomSearch, ctypes.c_uint64(seed),
                               ctypes.byref(i) if i is not None else None, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def onDemandTruncatedCostVolume(matchFunc, feature_vol_s, feature_vol_t, searchOffset, disp, radius=1, ctx=None):
    """OnDemand*CostVolume(source, target, searchSpace).truncatedCostVolume(disp, radius), correlation/on_demand_cost_volume.h:187-310:
    the costs around each pixel's disparity evaluated on demand -> (H,W,2r+1) or (H,W,2r+1,2r+1) float32."""
    ctx = _ctx_for(ctx, feature_vol_s)
    so, dims = _search_arg(searchOffset)
    n = 2 * radius + 1
    out = _empty(feature_vol_s, tuple(feature_vol_s.shape[:2]) + (n,) * dims, np.float32)
    s, t, d, o = describe(feature_vol_s), describe(feature_vol_t), describe(disp), describe(out)
    ctx.check(lib.stevi_on_demand_truncated_cost_volume(ctx.handle, matchFunc, ctypes.byref(s), ctypes.byref(t), dims, so, ctypes.byref(d),
                                                        radius, ctypes.byref(o)))
    return out


def refineFromCostVolume(cv, idx, h_radius, v_radius, kernel=PARABOLA, return_tcv=False, ctx=None):
    """truncatedCostVolume<Same>(radius 1) + refineDispCostInterpolation fused into one pass over the winners."""
    ctx = _ctx_for(ctx, cv)
    out = _empty(cv, cv.shape[:2], np.float32)
    tcv = _empty(cv, tuple(cv.shape[:2]) + (3,), np.float32) if return_tcv else None
    a, i, o = describe(cv), describe(idx), describe(out)
    t = describe(tcv) if tcv is not None else None
    ctx.check(lib.stevi_refine_from_cost_volume(ctx.handle, ctypes.byref(a), ctypes.byref(i), h_radius, v_radius, kernel,
                                                ctypes.byref(t) if t is not None else None, ctypes.byref(o)))
    return (tcv, out) if return_tcv else out


def averagePoolingDownsample(img, windows=2, ctx=None):
    """interpolation/downsampling.h:67-118 with DownSampleWindows(windows) or (horizontal, vertical)."""
    ctx = _ctx_for(ctx, img)
    wh, wv = (windows, windows) if np.isscalar(windows) else windows
    out = _empty(img, ((img.shape[0] + wv - 1) // wv, (img.shape[1] + wh - 1) // wh), np.float32)
    a, o = describe(img), describe(out)
    ctx.check(lib.stevi_avgpool_downsample(ctx.handle, ctypes.byref(a), wh, wv, ctypes.byref(o)))
    return out


def meanFilter2D(h_radius, v_radius, in_data, ctx=None):
    """correlation_base.h:881-1004: (H,W) or (H,W,C) u8|f32 -> (H,W) f32 box mean by the reference's running sums (interior
    bit-identical; the border it leaves uninitialised is 0). None when the window does not fit the image."""
    ctx = _ctx_for(ctx, in_data)
    out = _empty(in_data, in_data.shape[:2], np.float32)
    a, o = describe(in_data), describe(out)
    rc = lib.stevi_mean_filter_2d(ctx.handle, ctypes.byref(a), h_radius, v_radius, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def sigmaFilter(h_radius, v_radius, mean, in_data, ctx=None):
    """correlation_base.h:1006-1077: sqrt of the box sum of (in - mean(i,j))^2 on the interior, 0 on the border."""
    ctx = _ctx_for(ctx, in_data)
    out = _empty(in_data, in_data.shape[:2], np.float32)
    m, a, o = describe(mean), describe(in_data), describe(out)
    rc = lib.stevi_sigma_filter(ctx.handle, ctypes.byref(m), ctypes.byref(a), h_radius, v_radius, ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return out


def computeGuidedCV(matchFunc, img_l, img_r, h_radius, v_radius, guide, upscale_disp_radius=2, dDir=RIGHT_TO_LEFT, ctx=None):
    """hierarchical.h:73-233 on the images (unfold + features fused in). Returns (truncated_cost_volume, disp_estimate),
    or None where the reference returns an empty volume."""
    ctx = _ctx_for(ctx, img_l)
    ws = img_r.shape[1] if dDir == RIGHT_TO_LEFT else img_l.shape[1]
    tcv = _empty(img_l, (img_l.shape[0], ws, 2 * upscale_disp_radius + 1), np.float32)
    disp = _empty(img_l, (img_l.shape[0], ws), np.int32)
    l, r, g, t, d = describe(img_l), describe(img_r), describe(guide), describe(tcv), describe(disp)
    rc = lib.stevi_guided_cost_volume(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), h_radius, v_radius, ctypes.byref(g),
                                      upscale_disp_radius, dDir, ctypes.byref(t), ctypes.byref(d))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return tcv, disp


def computeGuidedCVFromFeatures(matchFunc, feature_vol_l, feature_vol_r, disp_guide, upscale_disp_radius=2, dDir=RIGHT_TO_LEFT, ctx=None):
    """computeGuidedCV with the reference's own operands, hierarchical.h:73-233: (H,W,F) f32 feature volumes as the caller prepared them
    and the (h_g,w_g) int32 guide. Returns (truncated_cost_volume, disp_estimate), or None where the reference returns an empty volume."""
    ctx = _ctx_for(ctx, feature_vol_l)
    ws = feature_vol_r.shape[1] if dDir == RIGHT_TO_LEFT else feature_vol_l.shape[1]
    tcv = _empty(feature_vol_l, (feature_vol_l.shape[0], ws, 2 * upscale_disp_radius + 1), np.float32)
    disp = _empty(feature_vol_l, (feature_vol_l.shape[0], ws), np.int32)
    l, r, g, t, d = describe(feature_vol_l), describe(feature_vol_r), describe(disp_guide), describe(tcv), describe(disp)
    rc = lib.stevi_guided_cost_volume_features(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), ctypes.byref(g), upscale_disp_radius, dDir,
                                               ctypes.byref(t), ctypes.byref(d))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return tcv, disp


def hiearchicalTruncatedCostVolume(matchFunc, depth, img_l, img_r, h_radius, v_radius, disp_width, upscale_disp_radius=2,
                                   dDir=RIGHT_TO_LEFT, ctx=None):
    """hierarchical.h:235-314 (sic). h_radius / v_radius: scalars, or depth+1 values coarsest level first.
    Returns (truncated_cost_volume, disp_estimate) like OffsetedCostVolume."""
    ctx = _ctx_for(ctx, img_l)
    tcv = _empty(img_l, tuple(img_l.shape) + (2 * upscale_disp_radius + 1,), np.float32)
    disp = _empty(img_l, img_l.shape, np.int32)
    arr = ctypes.c_int32 * (depth + 1)
    hrs = arr(*([int(h_radius)] * (depth + 1) if np.isscalar(h_radius) else [int(v) for v in h_radius]))
    vrs = arr(*([int(v_radius)] * (depth + 1) if np.isscalar(v_radius) else [int(v) for v in v_radius]))
    l, r, t, d = describe(img_l), describe(img_r), describe(tcv), describe(disp)
    rc = lib.stevi_hierarchical(ctx.handle, matchFunc, depth, ctypes.byref(l), ctypes.byref(r), hrs, vrs, 0, 0, disp_width,
                                upscale_disp_radius, dDir, ctypes.byref(t), ctypes.byref(d))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return tcv, disp


ROTATE0, ROTATE90, ROTATE180, ROTATE270 = range(4)


def _padding_arg(padding):
    if padding is None:
        return None
    return (ctypes.c_int32 * 4)(*[int(v) for v in padding])


def unfold(h_radius, v_radius, img, padding=None, orientation=ROTATE0, T_O=None, ctx=None):
    """unfold.h:248-344. padding = (left, top, right, bottom) or None for PaddingMargins() (automatic)."""
    ctx = _ctx_for(ctx, img)
    hs, vs = 2 * h_radius + 1, 2 * v_radius + 1
    pl, pt, pr, pb = (h_radius, v_radius, h_radius, v_radius) if padding is None else padding
    nch = img.shape[2] if len(img.shape) == 3 else 1
    np_in = _DT2NP[_torch_dtype_code(img)] if _is_torch(img) else img.dtype
    out = _empty(img, (img.shape[0] - vs + pt + pb + 1, img.shape[1] - hs + pl + pr + 1, hs * vs * nch), T_O or np_in)
    a, o = describe(img), describe(out)
    ctx.check(lib.stevi_unfold(ctx.handle, ctypes.byref(a), h_radius, v_radius, _padding_arg(padding), orientation, ctypes.byref(o)))
    return out


def compressor_geometry(mask):
    """(n_features, width, height, margins(left, top, right, bottom)) of UnFoldCompressor(mask), unfold.h:48-107."""
    mask = np.asarray(mask)
    ii, jj = np.nonzero(mask > 0)
    v_off, h_off = mask.shape[0] // 2, mask.shape[1] // 2
    min_h, max_h = min(0, int((ii - v_off).min())), max(0, int((ii - v_off).max()))
    min_w, max_w = min(0, int((jj - h_off).min())), max(0, int((jj - h_off).max()))
    return len(np.unique(mask[mask > 0])), max_w - min_w + 1, max_h - min_h + 1, (-min_w, -min_h, max_w, max_h)


def unfoldCompressed(mask, img, padding=None, ctx=None):
    """unfold(UnFoldCompressor(mask), img, padding), unfold.h:346-471. mask: 2-D int array (host)."""
    ctx = _ctx_for(ctx, img)
    mask = np.ascontiguousarray(mask, np.int32)
    nf, w, h, mg = compressor_geometry(mask)
    pl, pt, pr, pb = mg if padding is None else padding
    nch = img.shape[2] if len(img.shape) == 3 else 1
    out = _empty(img, (img.shape[0] - h + pt + pb + 1, img.shape[1] - w + pl + pr + 1, nf * nch), np.float32)
    m, a, o = describe(mask), describe(img), describe(out)
    ctx.check(lib.stevi_unfold_compressed(ctx.handle, ctypes.byref(m), ctypes.byref(a), _padding_arg(padding), ctypes.byref(o)))
    return out


def censusFeatures(feat, semantics=COMPAT, ctx=None):
    """census.h:69-115 on an (H,W,F) feature volume."""
    ctx = _ctx_for(ctx, feat)
    out = _empty(feat, tuple(feat.shape[:2]) + ((feat.shape[2] - 1) // 32 + 1,), np.uint32)
    a, o = describe(feat), describe(out)
    ctx.check(lib.stevi_census_features(ctx.handle, ctypes.byref(a), semantics, ctypes.byref(o)))
    return out


def getFeatureVolumeForMatchFunc(matchFunc, img, h_radius, v_radius, ctx=None):
    """cross_correlations.h:645-722 applied to unfold(h_radius, v_radius, img), float images."""
    ctx = _ctx_for(ctx, img)
    out = _empty(img, tuple(img.shape) + ((2 * h_radius + 1) * (2 * v_radius + 1),), np.float32)
    a, o = describe(img), describe(out)
    ctx.check(lib.stevi_feature_volume(ctx.handle, matchFunc, ctypes.byref(a), h_radius, v_radius, ctypes.byref(o)))
    return out


def localMatchingPipeline(matchFunc, img_l, img_r, h_radius, v_radius, disp_width, kernel=PARABOLA, dDir=RIGHT_TO_LEFT,
                          return_all=False, layout="ref", out=None, ctx=None):
    """unfoldBasedCostVolume -> extractSelectedIndex -> truncatedCostVolume(radius 1) -> refineDispCostInterpolation in one
    call. Returns the refined (H,W) f32 map, or (cv, idx, tcv, refined) with ``return_all``. ``out`` = preallocated
    refined map or (cv, idx, tcv, refined) tuple (entries may be None)."""
    ctx = _ctx_for(ctx, img_l)
    H = img_l.shape[0]
    ws = img_r.shape[1] if dDir == RIGHT_TO_LEFT else img_l.shape[1]
    cv = idx = tcv = None
    if out is not None:
        if isinstance(out, tuple):
            cv, idx, tcv, refined = out
        else:
            refined = out
    else:
        refined = _empty(img_l, (H, ws), np.float32)
        if return_all:
            cv = _empty(img_l, (H, ws, disp_width), np.float32, _volume_order(layout))
            idx = _empty(img_l, (H, ws), np.int32)
            tcv = _empty(img_l, (H, ws, 3), np.float32)
    l, r, o = describe(img_l), describe(img_r), describe(refined)
    d = [describe(x) if x is not None else None for x in (cv, idx, tcv)]
    ref = [ctypes.byref(x) if x is not None else None for x in d]
    rc = lib.stevi_local_matching_pipeline(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), h_radius, v_radius, disp_width,
                                           dDir, kernel, ref[0], ref[1], ref[2], ctypes.byref(o))
    if rc == E_SHAPE:
        return None
    ctx.check(rc)
    return (cv, idx, tcv, refined) if (return_all or isinstance(out, tuple)) else refined


class StBand(ctypes.Structure):
    _fields_ = [("img_row0", ctypes.c_int32), ("img_height", ctypes.c_int32), ("out_row0", ctypes.c_int32),
                ("guide_row0", ctypes.c_int32), ("guide_height", ctypes.c_int32)]


def computeGuidedCVBand(matchFunc, img_l, img_r, h_radius, v_radius, guide, img_row0, img_height, out_row0, n_rows,
                        guide_row0, guide_height, upscale_disp_radius=2, dDir=RIGHT_TO_LEFT, ctx=None):
    """Rows out_row0 .. out_row0+n_rows-1 of computeGuidedCV on a frame of img_height rows, from row bands of the images
    (starting at global row img_row0) and of the guide (starting at global guide row guide_row0 of guide_height)."""
    ctx = _ctx_for(ctx, img_l)
    ws = img_r.shape[1] if dDir == RIGHT_TO_LEFT else img_l.shape[1]
    tcv = _empty(img_l, (n_rows, ws, 2 * upscale_disp_radius + 1), np.float32)
    disp = _empty(img_l, (n_rows, ws), np.int32)
    band = StBand(int(img_row0), int(img_height), int(out_row0), int(guide_row0), int(guide_height))
    l, r, g, t, d = describe(img_l), describe(img_r), describe(guide), describe(tcv), describe(disp)
    ctx.check(lib.stevi_guided_cost_volume_band(ctx.handle, matchFunc, ctypes.byref(l), ctypes.byref(r), h_radius, v_radius,
                                                ctypes.byref(g), upscale_disp_radius, dDir, ctypes.byref(band),
                                                ctypes.byref(t), ctypes.byref(d)))
    return tcv, disp

"""TEST INFRASTRUCTURE — not product code.

ctypes/numpy front-end to the two CPU checkers:

* ``oracle.port``  — ``oracle/libstevi_oracle.so``: our plain-C restatement (``oracle/stevi_oracle.c``).
* ``oracle.ref``   — ``oracle/_ref/libstevi_ref.so``: the UNMODIFIED reference headers compiled from
  ``/root/reference`` by ``oracle/Makefile`` (``None`` when it was never built).
* ``oracle.ref_ibr`` — ``oracle/_ref/libstevi_ref_ibr.so``: the reference's ``correlation/image_based_refinement.h``,
  unmodified, with the Eigen calls inside it served by the stand-in of ``oracle/shims/Eigen`` + ``oracle/small_qr.h``.

Both expose the same method names. Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this module; ``libstevi_b200`` never does.
All arrays are dense C-order numpy arrays.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

# enums (matching_costs.h:38-53, correlation_base.h:31-45, cost_based_refinement.h:30-35)
CC, NCC, SSD, SAD, ZCC, ZNCC, ZSSD, ZSAD, HAMMING, CENSUS = 0, 1, 2, 3, 4, 5, 6, 7, 10, 11
COST, SCORE = 0, 1
L2R, R2L = 0, 1
EQUIANGULAR, PARABOLA, GAUSSIAN = 0, 1, 2
TCV_SAME, TCV_REVERSED, TCV_BOTH = 0, 1, 2
SCORE_FUNCS = (CC, NCC, ZCC, ZNCC)


def strategy_of(mf):
    return SCORE if mf in SCORE_FUNCS else COST


def build(ref=True):
    """(Re)build the checkers. Building the checker is not using it."""
    targets = ["oracle"] + (["ref"] if ref else [])
    subprocess.run(["make", "-s", "-C", _HERE] + targets, check=True)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


class _Checker:
    def __init__(self, path, prefix):
        self.lib = ctypes.CDLL(path)
        self.prefix = prefix
        self.path = path

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def has(self, name):
        return hasattr(self.lib, self.prefix + name)

    def set_num_threads(self, n):
        if self.prefix == "ref_":
            self.lib.ref_set_num_threads(int(n))

    # --- features -----------------------------------------------------------------
    def unfold(self, img, hr, vr):
        h, w = img.shape
        F = (2 * hr + 1) * (2 * vr + 1)
        if img.dtype == np.uint8:
            out = np.empty((h, w, F), np.uint8)
            self._f("unfold_u8")(_p(_c(img, np.uint8)), h, w, hr, vr, _p(out))
        else:
            out = np.empty((h, w, F), np.float32)
            self._f("unfold_f32")(_p(_c(img, np.float32)), h, w, hr, vr, _p(out))
        return out

    def unfold_general(self, img, hr, vr, padding=None, orientation=0):
        img = _c(img, np.float32)
        h, w = img.shape[:2]
        c = img.shape[2] if img.ndim == 3 else 0
        pad = np.asarray([-1, -1, -1, -1] if padding is None else padding, np.int32)
        oh, ow = ctypes.c_int(), ctypes.c_int()
        f = self._f("unfold_general_f32")
        f(_p(img), h, w, c, hr, vr, _p(pad), orientation, None, ctypes.byref(oh), ctypes.byref(ow))
        out = np.empty((oh.value, ow.value, (2 * hr + 1) * (2 * vr + 1) * max(c, 1)), np.float32)
        f(_p(img), h, w, c, hr, vr, _p(pad), orientation, _p(out), ctypes.byref(oh), ctypes.byref(ow))
        return out

    def unfold_compressed(self, mask, img, padding=None):
        img = _c(img, np.float32)
        mask = _c(mask, np.int32)
        h, w = img.shape[:2]
        c = img.shape[2] if img.ndim == 3 else 0
        pad = np.asarray([-1, -1, -1, -1] if padding is None else padding, np.int32)
        oh, ow, nf = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        f = self._f("unfold_compressed_f32")
        args = (_p(mask), mask.shape[0], mask.shape[1], _p(img), h, w, c, _p(pad))
        f(*args, None, ctypes.byref(oh), ctypes.byref(ow), ctypes.byref(nf))
        out = np.empty((oh.value, ow.value, nf.value * max(c, 1)), np.float32)
        f(*args, _p(out), ctypes.byref(oh), ctypes.byref(ow), ctypes.byref(nf))
        return out

    def census_features(self, fv):
        fv = _c(fv, np.float32)
        h, w, F = fv.shape
        out = np.zeros((h, w, (F - 1) // 32 + 1), np.uint32)
        self._f("census_features_f32")(_p(fv), h, w, F, _p(out))
        return out

    def census(self, img, hr, vr):
        h, w = img.shape
        F = (2 * hr + 1) * (2 * vr + 1)
        nw = (F - 1) // 32 + 1
        out = np.zeros((h, w, nw), np.uint32)
        if img.dtype == np.uint8:
            self._f("census_u8")(_p(_c(img, np.uint8)), h, w, hr, vr, _p(out))
        else:
            self._f("census_f32")(_p(_c(img, np.float32)), h, w, hr, vr, _p(out))
        return out

    def feature_volume(self, mf, img, hr, vr):
        h, w = img.shape
        out = np.empty((h, w, (2 * hr + 1) * (2 * vr + 1)), np.float32)
        rc = self._f("feature_volume_f32")(mf, _p(_c(img, np.float32)), h, w, hr, vr, _p(out))
        assert rc == 0
        return out

    # --- cost volumes ---------------------------------------------------------------
    def cost_volume(self, mf, left, right, hr, vr, D, direction=R2L):
        """unfoldBasedCostVolume; returns None where the reference returns an empty array."""
        h, wl = left.shape
        h2, wr = right.shape
        if h != h2:
            return None
        ws = wr if direction == R2L else wl
        if mf == CENSUS:
            out = np.empty((h, ws, D), np.uint16)
            if left.dtype == np.uint8:
                rc = self._f("census_cost_volume_u8")(_p(_c(left, np.uint8)), _p(_c(right, np.uint8)),
                                                      h, wl, wr, hr, vr, D, direction, _p(out))
            else:
                rc = self._f("census_cost_volume_f32")(_p(_c(left, np.float32)), _p(_c(right, np.float32)),
                                                       h, wl, wr, hr, vr, D, direction, _p(out))
        elif left.dtype == np.uint8 and mf in (CC, SSD, SAD) and self.has("cost_volume_u8"):
            # the reference's own uint8 instantiation (compiled reference only)
            out = np.empty((h, ws, D), np.float32)
            rc = self._f("cost_volume_u8")(mf, _p(_c(left, np.uint8)), _p(_c(right, np.uint8)),
                                           h, wl, wr, hr, vr, D, direction, _p(out))
        else:
            out = np.empty((h, ws, D), np.float32)
            rc = self._f("cost_volume_f32")(mf, _p(_c(left, np.float32)), _p(_c(right, np.float32)),
                                            h, wl, wr, hr, vr, D, direction, _p(out))
        return out if rc == 0 else None

    # --- feature-level forms (caller-supplied float feature volumes) --------------------
    def features_for_matchfunc(self, mf, fv):
        fv = _c(fv, np.float32)
        h, w, F = fv.shape
        out = np.empty((h, w, F), np.float32)
        rc = self._f("features_for_matchfunc_f32")(mf, _p(fv), h, w, F, _p(out))
        assert rc == 0
        return out

    def aggregate_cost(self, mf, fl, fr, lower, upper, direction=R2L):
        """aggregateCost (searchOffset<1>(lower, upper) overload) on processed features; None where the reference returns empty."""
        fl, fr = _c(fl, np.float32), _c(fr, np.float32)
        h, wl, F = fl.shape
        if fr.shape[0] != h or upper < lower:
            return None
        wr = fr.shape[1]
        out = np.empty((h, wr if direction == R2L else wl, upper - lower + 1), np.float32)
        rc = self._f("aggregate_cost_f32")(mf, _p(fl), _p(fr), h, wl, wr, F, lower, upper, direction, _p(out))
        return out if rc == 0 else None

    def feature_volume_to_cost_volume(self, mf, fl, fr, D, direction=R2L):
        if self.has("feature_volume_to_cost_volume_f32"):   # the compiled reference has the template itself
            fl, fr = _c(fl, np.float32), _c(fr, np.float32)
            h, wl, F = fl.shape
            wr = fr.shape[1]
            out = np.empty((h, wr if direction == R2L else wl, D), np.float32)
            rc = self._f("feature_volume_to_cost_volume_f32")(mf, _p(fl), _p(fr), h, wl, wr, F, D, direction, _p(out))
            return out if rc == 0 else None
        return self.aggregate_cost(mf, self.features_for_matchfunc(mf, fl), self.features_for_matchfunc(mf, fr), 0, D - 1, direction)

    def cost_volume_compressed(self, mf, mask, left, right, D, direction=R2L):
        if self.has("cost_volume_compressed_f32"):
            left, right, mask = _c(left, np.float32), _c(right, np.float32), _c(mask, np.int32)
            h, wl = left.shape[:2]
            wr = right.shape[1]
            c = left.shape[2] if left.ndim == 3 else 0
            out = np.empty((h, wr if direction == R2L else wl, D), np.float32)
            rc = self._f("cost_volume_compressed_f32")(mf, _p(mask), mask.shape[0], mask.shape[1], _p(left), _p(right), h, wl, wr, c,
                                                       D, direction, _p(out))
            return out if rc == 0 else None
        return self.feature_volume_to_cost_volume(mf, self.unfold_compressed(mask, left), self.unfold_compressed(mask, right), D, direction)

    def cost_volume_multichannel(self, mf, left, right, hr, vr, D, direction=R2L):
        if self.has("cost_volume_multichannel_f32"):
            left, right = _c(left, np.float32), _c(right, np.float32)
            h, wl, c = left.shape
            wr = right.shape[1]
            out = np.empty((h, wr if direction == R2L else wl, D), np.float32)
            rc = self._f("cost_volume_multichannel_f32")(mf, _p(left), _p(right), h, wl, wr, c, hr, vr, D, direction, _p(out))
            return out if rc == 0 else None
        return self.feature_volume_to_cost_volume(mf, self.unfold_general(left, hr, vr), self.unfold_general(right, hr, vr), D, direction)

    def cost_volume_2d(self, mf, left, right, hr, vr, search, direction=R2L):
        """unfoldBased2dDisparityCostVolume with searchOffset<2>(*search), search = (lower0, upper0, lower1, upper1);
        (h, w, dH, dW) float32, None where the reference returns an empty array."""
        if left.shape != right.shape:
            return None
        h, w = left.shape
        so = np.asarray(search, np.int32)
        dH, dW = int(so[1] - so[0] + 1), int(so[3] - so[2] + 1)
        if dH <= 0 or dW <= 0:
            return None
        out = np.empty((h, w, dH, dW), np.float32)
        rc = self._f("cost_volume_2d_f32")(mf, _p(_c(left, np.float32)), _p(_c(right, np.float32)), h, w, hr, vr, _p(so), direction, _p(out))
        return out if rc == 0 else None

    def wta_2d(self, vol, strategy):
        h, w, d1, d2 = vol.shape
        out = np.empty((h, w, 2), np.int32)
        self._f("wta_2d_f32")(_p(_c(vol, np.float32)), h, w, d1, d2, strategy, _p(out))
        return out

    def index2d_to_disp(self, idx, search):
        h, w, _ = idx.shape
        out = np.empty((h, w, 2), np.int32)
        self._f("index2d_to_disp")(_p(_c(idx, np.int32)), h, w, _p(np.asarray(search, np.int32)), _p(out))
        return out

    def sgm(self, cv, ndir, strategy, P1, P2, Pout=100.0, margins=(0, 0, 0, 0)):
        h, w, D = cv.shape
        out = np.empty((h, w, D), np.float32)
        m = np.asarray(margins, np.int32)
        f32 = ctypes.c_float
        # Q8: with 16 directions the reference's OpenMP loops race (lines of one direction share pixels); its
        # single-threaded result is the defined one
        serial = self.prefix == "ref_" and ndir == 16
        if serial:
            nthreads = self.lib.ref_num_threads()
            self.lib.ref_set_num_threads(1)
        try:
            return self._sgm(cv, ndir, strategy, P1, P2, Pout, m, out, f32, h, w, D)
        finally:
            if serial:
                self.lib.ref_set_num_threads(nthreads)

    def sgm_corrected(self, cv, ndir, strategy, P1, P2, Pout=100.0, margins=(0, 0, 0, 0)):
        """STEVI_CORRECTED semantics (the product's definition, not the reference's behaviour; port only)."""
        cv = _c(cv, np.float32)
        h, w, D = cv.shape
        out = np.empty((h, w, D), np.float32)
        m = np.asarray(margins, np.int32)
        f32 = ctypes.c_float
        rc = self._f("sgm_f32_ex")(_p(cv), h, w, D, ndir, strategy, f32(P1), f32(P2), f32(Pout), _p(m), 1, _p(out))
        assert rc == 0
        return out

    def _sgm(self, cv, ndir, strategy, P1, P2, Pout, m, out, f32, h, w, D):
        if cv.dtype == np.uint16:
            rc = self._f("sgm_u16")(_p(_c(cv, np.uint16)), h, w, D, ndir, strategy, f32(P1), f32(P2), f32(Pout), _p(m), _p(out))
        else:
            rc = self._f("sgm_f32")(_p(_c(cv, np.float32)), h, w, D, ndir, strategy, f32(P1), f32(P2), f32(Pout), _p(m), _p(out))
        assert rc == 0
        return out

    def wta(self, vol, strategy):
        h, w, D = vol.shape
        out = np.empty((h, w), np.int32)
        if vol.dtype == np.uint16:
            self._f("wta_u16")(_p(_c(vol, np.uint16)), h, w, D, strategy, _p(out))
        else:
            self._f("wta_f32")(_p(_c(vol, np.float32)), h, w, D, strategy, _p(out))
        return out

    def index_to_disp(self, idx, direction=R2L, offset=0):
        h, w = idx.shape
        out = np.empty((h, w), np.int32)
        self._f("index_to_disp")(_p(_c(idx, np.int32)), h, w, direction, offset, _p(out))
        return out

    def selected_cost(self, cv, idx):
        h, w, D = cv.shape
        out = np.empty((h, w), np.float32)
        self._f("selected_cost_f32")(_p(_c(cv, np.float32)), _p(_c(idx, np.int32)), h, w, D, _p(out))
        return out

    def truncated_cost_volume(self, cv, idx, hr, vr, radius, direction=R2L, tdir=TCV_SAME):
        h, w, D = cv.shape
        T = 4 * radius + 1 if tdir == TCV_BOTH else 2 * radius + 1
        out = np.empty((h, w, T), np.float32)
        self._f("truncated_cost_volume_f32")(_p(_c(cv, np.float32)), _p(_c(idx, np.int32)), h, w, D,
                                             hr, vr, radius, direction, tdir, _p(out))
        return out

    def in_bound_domain(self, idx, width, hr, vr, radius, direction=R2L, tdir=TCV_SAME):
        h, w = idx.shape
        T = 4 * radius + 1 if tdir == TCV_BOTH else 2 * radius + 1
        out = np.empty((h, w, T), np.float32)
        self._f("in_bound_domain_f32")(_p(_c(idx, np.int32)), h, w, width, hr, vr, radius, direction, tdir, _p(out))
        return out

    def refine(self, tcv, idx, kernel=PARABOLA):
        h, w, T = tcv.shape
        out = np.empty((h, w), np.float32)
        rc = self._f("refine_disp_cost_interp")(_p(_c(tcv, np.float32)), _p(_c(idx, np.int32)), h, w, T, kernel, _p(out))
        return out if rc == 0 else None

    def truncated_cost_volume_2d(self, cv, idx, radius0=255, radius1=255):
        """truncatedBidirectionaCostVolume; the radii are the reference's uint8_t arguments (default uint8_t(-1) = 255)."""
        h, w, d1, d2 = cv.shape
        r0 = max(radius0 & 255, 1)
        r1 = (radius1 & 255) or r0
        out = np.empty((h, w, 2 * r0 + 1, 2 * r1 + 1), np.float32)
        self._f("truncated_cost_volume_2d_f32")(_p(_c(cv, np.float32)), _p(_c(idx, np.int32)), h, w, d1, d2, radius0, radius1, _p(out))
        return out

    def refine_2d(self, tcv, raw, kernel=PARABOLA, anisotropic=False):
        """refineDisp2dCostInterpolation; None where the reference returns an empty array."""
        h, w, n0, n1 = tcv.shape
        out = np.empty((h, w, 2), np.float32)
        rc = self._f("refine_disp_2d_cost_interp")(_p(_c(tcv, np.float32)), _p(_c(raw, np.int32)), h, w, n0, n1, kernel, int(anisotropic), _p(out))
        return out if rc == 0 else None

    def refine_2d_patch(self, tcv, raw, kernel=PARABOLA):
        """refineDisp2dCostPatchInterpolation (``ref_ibr`` runs the reference's code on the Eigen stand-in's Gauss-Jordan inverse)."""
        h, w, n0, n1 = tcv.shape
        out = np.empty((h, w, 2), np.float32)
        rc = self._f("refine_disp_2d_cost_patch_interp")(_p(_c(tcv, np.float32)), _p(_c(raw, np.int32)), h, w, n0, n1, kernel, _p(out))
        return out if rc == 0 else None

    def variable_cost_volume_aggregation(self, cv, windows):
        """variableCostVolumeAggregation<float, float>; None where the reference returns an empty array."""
        h, w, d = cv.shape
        hw, ww, n0, n1 = windows.shape
        out = np.empty((h, w, d), np.float32)
        rc = self._f("variable_cost_volume_aggregation_f32")(_p(_c(cv, np.float32)), h, w, d, _p(_c(windows, np.float32)), hw, ww, n0, n1, _p(out))
        return out if rc == 0 else None

    def avgpool2(self, img):
        h, w = img.shape
        out = np.empty(((h + 1) // 2, (w + 1) // 2), np.float32)
        self._f("avgpool2_f32")(_p(_c(img, np.float32)), h, w, _p(out))
        return out

    def guided_cv(self, mf, left, right, hr, vr, guide, radius=2, direction=R2L):
        h, wl = left.shape
        wr = right.shape[1]
        ws = wr if direction == R2L else wl
        tcv = np.empty((h, ws, 2 * radius + 1), np.float32)
        disp = np.empty((h, ws), np.int32)
        g = _c(guide, np.int32)
        rc = self._f("guided_cv_f32")(mf, _p(_c(left, np.float32)), _p(_c(right, np.float32)), h, wl, wr, hr, vr,
                                      _p(g), g.shape[0], g.shape[1], radius, direction, _p(tcv), _p(disp))
        assert rc == 0
        return tcv, disp

    def guided_cv_features(self, mf, fl, fr, guide, radius=2, direction=R2L):
        fl, fr, g = _c(fl, np.float32), _c(fr, np.float32), _c(guide, np.int32)
        ws = fr.shape[1] if direction == R2L else fl.shape[1]
        tcv = np.empty((fl.shape[0], ws, 2 * radius + 1), np.float32)
        disp = np.empty((fl.shape[0], ws), np.int32)
        rc = self._f("guided_cv_features_f32")(mf, _p(fl), _p(fr), fl.shape[0], fl.shape[1], fr.shape[1], fl.shape[2], _p(g), g.shape[0], g.shape[1],
                                               radius, direction, _p(tcv), _p(disp))
        return (tcv, disp) if rc == 0 else None

    def hierarchical(self, mf, depth, left, right, hr, vr, D, radius=2, direction=R2L):
        h, w = left.shape
        tcv = np.empty((h, w, 2 * radius + 1), np.float32)
        disp = np.empty((h, w), np.int32)
        rc = self._f("hierarchical_f32")(mf, depth, _p(_c(left, np.float32)), _p(_c(right, np.float32)), h, w,
                                         hr, vr, D, radius, direction, _p(tcv), _p(disp))
        return (tcv, disp) if rc == 0 else None

    # --- correlation_base.h:881-1077 ------------------------------------------------------
    def mean_filter_2d(self, img, hr, vr, fill=np.nan):
        """meanFilter2D on (H,W) or (H,W,C) u8|f32. Only the interior [vr,H-vr) x [hr,W-hr) is defined by the reference; the
        rest of the returned array holds ``fill``. None if the window does not fit."""
        h, w = img.shape[:2]
        c = img.shape[2] if img.ndim == 3 else 0
        out = np.full((h, w), fill, np.float32)
        if img.dtype == np.uint8:
            rc = self._f("mean_filter_2d_u8")(_p(_c(img, np.uint8)), h, w, c, hr, vr, _p(out))
        else:
            rc = self._f("mean_filter_2d_f32")(_p(_c(img, np.float32)), h, w, c, hr, vr, _p(out))
        return out if rc == 0 else None

    def sigma_filter(self, mean, img, hr, vr, fill=np.nan):
        """sigmaFilter(h_radius, v_radius, mean, in_data); interior only, like mean_filter_2d."""
        h, w = img.shape[:2]
        c = img.shape[2] if img.ndim == 3 else 0
        out = np.full((h, w), fill, np.float32)
        if img.dtype == np.uint8:
            rc = self._f("sigma_filter_u8")(_p(_c(mean, np.float32)), _p(_c(img, np.uint8)), h, w, c, hr, vr, _p(out))
        else:
            rc = self._f("sigma_filter_f32")(_p(_c(mean, np.float32)), _p(_c(img, np.float32)), h, w, c, hr, vr, _p(out))
        return out if rc == 0 else None


    # --- feature-domain refinement (image_based_refinement.h) --------------------------
    def refine_barycentric(self, mf, fl, fr, idx, direction=R2L):
        fl, fr, idx = _c(fl, np.float32), _c(fr, np.float32), _c(idx, np.int32)
        out = np.empty(idx.shape, np.float32)
        rc = self._f("refine_barycentric_disp")(mf, _p(fl), _p(fr), fl.shape[0], fl.shape[1], fr.shape[1], fl.shape[2], _p(idx), direction, _p(out))
        return out if rc == 0 else None

    def refine_barycentric_symmetric(self, mf, radius, fl, fr, idx, width, direction=R2L):
        fl, fr, idx = _c(fl, np.float32), _c(fr, np.float32), _c(idx, np.int32)
        out = np.empty(idx.shape, np.float32)
        rc = self._f("refine_barycentric_symmetric_disp")(mf, radius, _p(fl), _p(fr), fl.shape[0], fl.shape[1], fr.shape[1], fl.shape[2],
                                                          _p(idx), width, direction, _p(out))
        return out if rc == 0 else None

    def refine_barycentric_2d(self, mf, contiguity, symmetric, fl, fr, idx, search, direction=R2L):
        fl, fr, idx = _c(fl, np.float32), _c(fr, np.float32), _c(idx, np.int32)
        so = np.asarray(search, np.int32)
        out = np.empty(idx.shape, np.float32)
        rc = self._f("refine_barycentric_2d_disp")(mf, contiguity, int(bool(symmetric)), _p(fl), _p(fr), fl.shape[0], fl.shape[1], fl.shape[2],
                                                   _p(idx), _p(so), direction, _p(out))
        return out if rc == 0 else None

    def refine_cost_symmetric(self, mf, fl, fr, idx, cv, direction=R2L):
        fl, fr, idx, cv = _c(fl, np.float32), _c(fr, np.float32), _c(idx, np.int32), _c(cv, np.float32)
        out = np.empty(idx.shape, np.float32)
        rc = self._f("refine_cost_symmetric_disp")(mf, _p(fl), _p(fr), fl.shape[0], fl.shape[1], fr.shape[1], fl.shape[2], _p(idx), _p(cv),
                                                   cv.shape[2], direction, _p(out))
        return out if rc == 0 else None

    # --- on-demand cost volumes and PatchMatch (on_demand_cost_volume.h, patchmatch.h) --
    def patch_match(self, mf, fs, ft, search, n_iter=5, n_random=4, init=None):
        fs, ft = _c(fs, np.float32), _c(ft, np.float32)
        so = np.asarray(search, np.int32)
        dims = len(so) // 2
        out = np.zeros(fs.shape[:2] + (dims,), np.int32)
        ini = _c(init, np.int32) if init is not None else None
        rc = self._f("patch_match")(mf, dims, _p(fs), _p(ft), fs.shape[0], fs.shape[1], ft.shape[0], ft.shape[1], fs.shape[2], _p(so), n_iter, n_random,
                                    _p(ini) if ini is not None else None, _p(out))
        return out if rc == 0 else None

    def on_demand_truncated_cost_volume(self, mf, fs, ft, search, disp, radius=1):
        fs, ft, disp = _c(fs, np.float32), _c(ft, np.float32), _c(disp, np.int32)
        so = np.asarray(search, np.int32)
        dims = len(so) // 2
        n = 2 * radius + 1
        out = np.zeros(fs.shape[:2] + (n,) * dims, np.float32)
        rc = self._f("on_demand_truncated_cost_volume")(mf, dims, _p(fs), _p(ft), fs.shape[0], fs.shape[1], ft.shape[0], ft.shape[1], fs.shape[2],
                                                        _p(so), _p(disp), radius, _p(out))
        return out if rc == 0 else None

    def refined_barycentric_disp(self, mf, left, right, hr, vr, width, prenormalize=False):
        left, right = _c(left, np.float32), _c(right, np.float32)
        out = np.empty(right.shape, np.float32)
        rc = self._f("refined_barycentric_disp_f32")(mf, _p(left), _p(right), left.shape[0], left.shape[1], right.shape[1], hr, vr, width,
                                                     int(bool(prenormalize)), _p(out))
        return out if rc == 0 else None


def _load():
    port_path = os.path.join(_HERE, "libstevi_oracle.so")
    ref_path = os.path.join(_HERE, "_ref", "libstevi_ref.so")
    if not os.path.exists(port_path):
        build(ref=False)
    p = _Checker(port_path, "orc_")
    p.lib.orc_refine_cost_triplet.restype = ctypes.c_float
    p.lib.orc_round_through_float.restype = ctypes.c_uint32
    r = _Checker(ref_path, "ref_") if os.path.exists(ref_path) else None
    # correlation/image_based_refinement.h compiled as its own library (oracle/ref_wrap_ibr.cpp): the reference's code with the
    # stand-in linear algebra of oracle/shims/Eigen
    ibr_path = os.path.join(_HERE, "_ref", "libstevi_ref_ibr.so")
    i = _Checker(ibr_path, "ref_") if os.path.exists(ibr_path) else None
    return p, r, i


port, ref, ref_ibr = _load()

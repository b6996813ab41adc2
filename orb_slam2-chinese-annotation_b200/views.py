"""ctypes mirrors of the plain view structs of include/orb_b200.h, shared by the product wrapper
(matcher.py) and the test-side oracle binding (the structs are interface types, not behaviour)."""
import ctypes as C
import numpy as np

FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48   # reference include/Frame.h:39-40


class GridParams(C.Structure):
    _fields_ = [("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
                ("inv_w", C.c_float), ("inv_h", C.c_float)]


class FrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("x", C.c_void_p), ("y", C.c_void_p), ("octave", C.c_void_p), ("angle", C.c_void_p),
                ("u_right", C.c_void_p), ("desc", C.c_void_p), ("grid", GridParams), ("scale_factors", C.c_void_p),
                ("n_levels", C.c_int32)]


class MapPointView(C.Structure):
    _fields_ = [("n", C.c_int32), ("valid", C.c_void_p), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p),
                ("proj_xr", C.c_void_p), ("level", C.c_void_p), ("view_cos", C.c_void_p), ("desc", C.c_void_p),
                ("blocks", C.c_void_p)]


class LastFrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("valid", C.c_void_p), ("xc", C.c_void_p), ("yc", C.c_void_p), ("zc", C.c_void_p),
                ("octave", C.c_void_p), ("angle", C.c_void_p), ("desc", C.c_void_p), ("blocks", C.c_void_p)]


class FrameBatchView(C.Structure):
    _fields_ = [("pairs", C.c_int32), ("stride", C.c_int32), ("n", C.c_void_p), ("x", C.c_void_p), ("y", C.c_void_p), ("octave", C.c_void_p),
                ("angle", C.c_void_p), ("u_right", C.c_void_p), ("desc", C.c_void_p), ("grid", GridParams), ("scale_factors", C.c_void_p),
                ("n_levels", C.c_int32)]


class MapPointBatchView(C.Structure):
    _fields_ = [("pairs", C.c_int32), ("stride", C.c_int32), ("n", C.c_void_p), ("valid", C.c_void_p), ("proj_x", C.c_void_p),
                ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p), ("level", C.c_void_p), ("view_cos", C.c_void_p), ("desc", C.c_void_p),
                ("blocks", C.c_void_p)]


class LastFrameBatchView(C.Structure):
    _fields_ = [("pairs", C.c_int32), ("stride", C.c_int32), ("n", C.c_void_p), ("valid", C.c_void_p), ("xc", C.c_void_p), ("yc", C.c_void_p),
                ("zc", C.c_void_p), ("octave", C.c_void_p), ("angle", C.c_void_p), ("desc", C.c_void_p), ("blocks", C.c_void_p)]


class Camera(C.Structure):
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("mbf", C.c_float)]


class ProjectedView(C.Structure):
    _fields_ = [("n", C.c_int32), ("valid", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p), ("ur", C.c_void_p),
                ("radius", C.c_void_p), ("min_level", C.c_void_p), ("max_level", C.c_void_p), ("desc", C.c_void_p),
                ("angle", C.c_void_p)]


class FeatVecView(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("node_ids", C.c_void_p), ("start", C.c_void_p), ("items", C.c_void_p)]


def _addr(a):
    return None if a is None else a.ctypes.data


class Frame:
    """The slice of ORB_SLAM2::Frame / KeyFrame the matcher reads: mvKeysUn (x, y, octave, angle),
    mvuRight, mDescriptors, image bounds and grid constants, mvScaleFactors."""

    def __init__(self, x, y, octave, angle, desc, scale_factors, bounds, u_right=None):
        self.x = np.ascontiguousarray(x, np.float32); self.y = np.ascontiguousarray(y, np.float32)
        self.octave = np.ascontiguousarray(octave, np.int32); self.angle = np.ascontiguousarray(angle, np.float32)
        self.desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        self.sf = np.ascontiguousarray(scale_factors, np.float32)
        self.N = len(self.x)
        mnx, mny, mxx, mxy = (np.float32(v) for v in bounds)
        self.bounds = (mnx, mny, mxx, mxy)
        # reference src/Frame.cc:164-165
        self.inv_w = np.float32(FRAME_GRID_COLS) / np.float32(mxx - mnx)
        self.inv_h = np.float32(FRAME_GRID_ROWS) / np.float32(mxy - mny)

    def view(self):
        g = GridParams(self.bounds[0], self.bounds[1], self.bounds[2], self.bounds[3], self.inv_w, self.inv_h)
        return FrameView(self.N, _addr(self.x), _addr(self.y), _addr(self.octave), _addr(self.angle), _addr(self.u_right),
                         _addr(self.desc), g, _addr(self.sf), len(self.sf))


class FeatureVector:
    """DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) flattened in ascending node id."""

    def __init__(self, mapping):
        ids = sorted(mapping)
        self.node_ids = np.array(ids, np.uint32)
        lens = [len(mapping[k]) for k in ids]
        self.start = np.zeros(len(ids) + 1, np.int32)
        self.start[1:] = np.cumsum(lens)
        self.items = np.array([i for k in ids for i in mapping[k]], np.uint32)

    def view(self):
        return FeatVecView(len(self.node_ids), _addr(self.node_ids), _addr(self.start), _addr(self.items))


def u8(a):
    return np.ascontiguousarray(np.asarray(a).astype(np.uint8))

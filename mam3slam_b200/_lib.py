"""Loader of libmam3slam_b200.so (CUDA kernels + C ABI, include/orbx.h + include/orbm.h).

The library is built in-tree by `__graft_entry__.build()` / `make -C mam3slam_b200/csrc`.  There is no CPU fallback:
if the library is missing, or no CUDA device is present when a compute entry point is called, this fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libmam3slam_b200.so")
_LIB = None


class OrbxConfig(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int),
                ("ini_th_fast", C.c_int), ("min_th_fast", C.c_int), ("device", C.c_int), ("max_batch", C.c_int)]


class NativeError(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(SO_PATH):
        raise NativeError(f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(SO_PATH)
    vp, i32, f32, u64 = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    ip = C.POINTER(C.c_int)
    L.orbx_device_count.restype = i32
    L.orbx_device_pci_bus_id.restype = i32
    L.orbx_device_pci_bus_id.argtypes = [i32, C.c_char_p, i32]
    L.orbx_pin_host.restype = i32
    L.orbx_pin_host.argtypes = [C.c_void_p, C.c_size_t]
    L.orbx_unpin_host.restype = i32
    L.orbx_unpin_host.argtypes = [C.c_void_p]
    L.orbx_strerror.restype = C.c_char_p
    L.orbx_strerror.argtypes = [i32]
    L.orbx_last_error.restype = C.c_char_p
    L.orbx_last_error.argtypes = [vp]
    L.orbx_create.argtypes = [C.POINTER(OrbxConfig), C.POINTER(vp)]
    L.orbx_destroy.argtypes = [vp]
    L.orbx_destroy.restype = None
    L.orbx_get_tables.argtypes = [vp, vp, vp, vp, vp, vp]
    L.orbx_max_keypoints.argtypes = [vp]
    L.orbx_extract.argtypes = [vp, vp, i32, i32, i32, i32, i32, vp, vp, i32, ip, ip]
    L.orbx_extract_batch.argtypes = [vp, vp, i32, i32, i32, i32, i32, i32, vp, vp, i32, vp, vp]
    L.orbx_extract_device.argtypes = [vp, vp, u64, i32, i32, i32, i32, i32, i32, vp]
    L.orbx_synchronize.argtypes = [vp]
    L.orbx_results_device.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(vp), ip, ip]
    L.orbx_results_batch_device.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), ip]
    L.orbx_fetch_results.argtypes = [vp, i32, vp, vp, i32, ip, ip]
    L.orbx_level_size.argtypes = [vp, i32, ip, ip]
    L.orbx_stereo_matches.argtypes = [vp, vp, f32, f32, vp, vp, i32, ip, ip]
    L.orbx_download_level.argtypes = [vp, i32, i32, i32, vp, i32]
    L.orbx_download_blurred.argtypes = [vp, i32, i32, vp, i32]
    L.orbx_download_candidates.argtypes = [vp, i32, i32, vp, i32, ip]
    L.orbx_download_level_keypoints.argtypes = [vp, i32, i32, vp, i32, ip]
    L.orbx_debug_sort_nodes.argtypes = [vp, vp, vp, i32, vp, i32]
    L.orbx_debug_distribute.argtypes = [vp, vp, i32, i32, i32, i32, i32, i32, vp, i32, ip]
    L.orbx_debug_distribute2.argtypes = [vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, i32, ip, ip]
    L.orbx_launch_count.restype = C.c_longlong
    L.orbx_launch_count.argtypes = [vp]
    L.orbx_stream.restype = vp
    L.orbx_stream.argtypes = [vp]
    L.orbx_set_stage_timing.argtypes = [vp, i32]
    L.orbx_stage_times.argtypes = [vp, vp]
    _bind_matcher(L)
    _LIB = L
    return L


class OrbmFrameView(C.Structure):
    _fields_ = [("n", C.c_int), ("kps", C.c_void_p), ("desc", C.c_void_p), ("min_x", C.c_float), ("min_y", C.c_float),
                ("max_x", C.c_float), ("max_y", C.c_float), ("on_device", C.c_int)]


class OrbmVocabView(C.Structure):
    _fields_ = [("nnodes", C.c_int), ("levels", C.c_int), ("child_off", C.c_void_p), ("children", C.c_void_p),
                ("node_desc", C.c_void_p), ("word_id", C.c_void_p), ("weight", C.c_void_p), ("weighting", C.c_int),
                ("norm", C.c_int)]


class OrbmFeatVec(C.Structure):
    _fields_ = [("n_nodes", C.c_int), ("node", C.c_void_p), ("off", C.c_void_p), ("idx", C.c_void_p)]


def _bind_matcher(L):
    vp, i32, f32, ll = C.c_void_p, C.c_int, C.c_float, C.c_longlong
    ip = C.POINTER(C.c_int)
    fvp, fwp = C.POINTER(OrbmFeatVec), C.POINTER(OrbmFrameView)
    L.orbm_strerror.restype = C.c_char_p
    L.orbm_strerror.argtypes = [i32]
    L.orbm_last_error.restype = C.c_char_p
    L.orbm_last_error.argtypes = [vp]
    L.orbm_create.argtypes = [i32, C.POINTER(vp)]
    L.orbm_destroy.argtypes = [vp]
    L.orbm_destroy.restype = None
    L.orbm_launch_count.restype = ll
    L.orbm_launch_count.argtypes = [vp]
    L.orbm_stream.restype = vp
    L.orbm_stream.argtypes = [vp]
    L.orbm_last_evaluations.restype = ll
    L.orbm_last_evaluations.argtypes = [vp]
    L.orbm_distance.argtypes = [vp, vp]
    L.orbm_frame_create.argtypes = [vp, fwp, C.POINTER(vp)]
    L.orbm_frame_update.argtypes = [vp, vp, fwp]
    L.orbm_frame_destroy.argtypes = [vp, vp]
    L.orbm_frame_destroy.restype = None
    L.orbm_frame_grid.argtypes = [vp, vp, vp, vp]
    L.orbm_features_in_area.argtypes = [vp, vp, f32, f32, f32, i32, i32, vp, i32, ip]
    L.orbm_search_local_map.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, f32, f32, vp, ip, vp]
    L.orbm_search_frame.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, f32, i32, i32, vp, ip, vp]
    L.orbm_search_sim3.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp, vp, vp, vp, i32, f32, vp, ip]
    L.orbm_search_initialization.argtypes = [vp, vp, vp, vp, i32, f32, i32, vp, ip]
    L.orbm_knn_merge_device.argtypes = [vp, vp, vp, C.c_size_t, i32, i32, i32, vp, vp, vp]
    L.orbm_fuse_candidates.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp, vp, vp, vp, f32, i32, vp, vp]
    L.orbm_search_by_sim3.argtypes = [vp, vp, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, f32, vp, ip]
    L.orbm_search_bow_kf_frame.argtypes = [vp, vp, fvp, vp, vp, fvp, f32, i32, vp, ip]
    L.orbm_search_bow_kf_kf.argtypes = [vp, vp, fvp, vp, vp, fvp, vp, f32, i32, vp, ip]
    L.orbm_search_triangulation.argtypes = [vp, vp, fvp, vp, vp, fvp, vp, vp, vp, vp, vp, i32, i32, i32, vp, ip]
    L.orbm_db_create.argtypes = [vp, vp, ll, i32, ll, C.POINTER(vp)]
    L.orbm_db_destroy.argtypes = [vp, vp]
    L.orbm_db_destroy.restype = None
    L.orbm_knn.argtypes = [vp, vp, vp, i32, i32, vp, vp]
    L.orbm_knn_device.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp]
    L.orbm_knn_device_tc.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp]
    L.orbm_xchg_create.argtypes = [vp, i32, i32, i32, i32, C.POINTER(vp)]
    L.orbm_xchg_ipc_handle.argtypes = [vp, vp]
    L.orbm_xchg_open.argtypes = [vp, i32, vp]
    L.orbm_xchg_open_local.argtypes = [vp, i32, vp]
    L.orbm_xchg_status.argtypes = [vp]
    L.orbm_xchg_destroy.argtypes = [vp]
    L.orbm_xchg_destroy.restype = None
    L.orbm_knn_sharded.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp, vp]
    L.orbm_knn_device_static.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp]
    L.orbm_debug_alloc_count.restype = C.c_longlong
    L.orbm_debug_alloc_count.argtypes = []
    L.orbm_knn_merge.argtypes = [vp, vp, i32, i32, i32, vp, vp]
    L.orbm_distinctive_descriptors.argtypes = [vp, vp, vp, i32, vp]
    L.orbm_track_batch_device.argtypes = [vp, i32, vp, vp, vp, i32, i32, vp, vp, i32, f32, i32, i32, vp, vp, vp]
    L.orbm_undistort_points.argtypes = [vp, vp, i32, vp, vp, i32, vp]
    L.orbm_undistort_keypoints_device.argtypes = [vp, i32, vp, vp, i32, i32, vp, vp, i32, vp, vp]
    L.orbm_vocab_create.argtypes = [vp, C.POINTER(OrbmVocabView), C.POINTER(vp)]
    L.orbm_vocab_destroy.argtypes = [vp, vp]
    L.orbm_vocab_destroy.restype = None
    L.orbm_bow_transform.argtypes = [vp, vp, vp, i32, i32, vp, vp, ip, vp, vp, ip, vp, vp, vp]
    L.orbm_bow_transform_device.argtypes = [vp, vp, i32, vp, vp, i32, i32, i32, vp, vp, vp]
    L.orbm_track_evaluations.restype = ll
    L.orbm_track_evaluations.argtypes = [vp]


def check_m(status, handle=None, what=""):
    if status == 0:
        return
    L = lib()
    msg = L.orbm_strerror(status).decode()
    if handle:
        detail = L.orbm_last_error(handle).decode()
        if detail:
            msg += ": " + detail
    raise NativeError(f"{what} failed ({status}): {msg}")


def check(status, handle=None, what=""):
    if status == 0:
        return
    L = lib()
    msg = L.orbx_strerror(status).decode()
    if handle:
        detail = L.orbx_last_error(handle).decode()
        if detail:
            msg += ": " + detail
    raise NativeError(f"{what} failed ({status}): {msg}")


def device_count():
    """Number of CUDA devices the library sees (0 on a CPU-only host)."""
    return int(lib().orbx_device_count())


def device_pci_bus_id(device):
    """PCI bus id of a CUDA device, lower case ("0000:1b:00.0")."""
    buf = C.create_string_buffer(32)
    check(lib().orbx_device_pci_bus_id(int(device), buf, 32), None, "orbx_device_pci_bus_id")
    return buf.value.decode().lower()

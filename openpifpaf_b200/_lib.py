"""ctypes binding of libpifpaf_b200.so (include/pifpaf_b200.h).

There is no CPU fallback: if the library is missing or fails to load, importing
the decoder raises.  The library is built in-tree by ``openpifpaf_b200/build.py``
(``__graft_entry__.build()``).
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'libpifpaf_b200.so')

PP_MAX_F, PP_MAX_L = 32, 64
PP_N_STAGES = 7
PP_OK, PP_ERR_ARG, PP_ERR_CAPACITY, PP_ERR_CUDA, PP_ERR_WORKSPACE, PP_ERR_STATIC_LIMIT = 0, -1, -2, -3, -4, -5
STATUS_BITS = {1: 'seed_cap', 2: 'caf_cap', 4: 'ann_cap', 8: 'src_cap'}
PP_STATUS_OVERFLOW_MASK = 15
PP_STATUS_SCORE_ABOVE_ONE = 16    # warning, not an error: see include/pifpaf_b200.h


class Shape(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ('B', 'F', 'L', 'H', 'W', 'stride')]


class Caps(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ('seed_cap', 'caf_cap', 'ann_cap', 'src_cap')]


class Initial(C.Structure):
    _fields_ = [('data', C.c_void_p), ('scales', C.c_void_p), ('n', C.c_void_p), ('cap', C.c_int32)]


class ImageMeta(C.Structure):
    _fields_ = [('offset_x', C.c_double), ('offset_y', C.c_double),
                ('scale_x', C.c_double), ('scale_y', C.c_double), ('flip_w1', C.c_double),
                ('rot_cos', C.c_float), ('rot_sin', C.c_float),
                ('rot_cx', C.c_float), ('rot_cy', C.c_float),
                ('rotate', C.c_int32), ('hflip', C.c_int32)]


class Inverse(C.Structure):
    _fields_ = [('meta', C.c_void_p), ('hflip_target', C.c_void_p)]


class FieldSet(C.Structure):
    _fields_ = [('cif', C.c_void_p), ('caf', C.c_void_p), ('H', C.c_int32), ('W', C.c_int32),
                ('stride', C.c_int32), ('cif_min_scale', C.c_float), ('caf_min_distance', C.c_float),
                ('caf_max_distance', C.c_float), ('cif_rest', C.c_void_p), ('caf_rest', C.c_void_p)]


class DetParams(C.Structure):
    _fields_ = [('cif_threshold', C.c_float), ('seed_threshold', C.c_float),
                ('seed_score_scale', C.c_float), ('cif_min_scale', C.c_float),
                ('nms_suppression', C.c_float), ('nms_suppression_soft', C.c_float),
                ('nms_instance_threshold', C.c_float), ('nms_iou_threshold', C.c_float),
                ('nms_iou_threshold_soft', C.c_float), ('nms', C.c_int32)]


class DetOutputs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        'hr', 'seeds', 'n_seeds', 'pre_field', 'pre_score', 'pre_box', 'n_pre',
        'det_field', 'det_score', 'det_box', 'det_source', 'n_det', 'status')]


class Params(C.Structure):
    _fields_ = [
        ('nms_instance_threshold', C.c_double),
        ('cif_threshold', C.c_float), ('seed_threshold', C.c_float),
        ('seed_score_scale', C.c_float), ('caf_threshold', C.c_float),
        ('cif_floor', C.c_float), ('cif_floor_compl', C.c_float),
        ('keypoint_threshold', C.c_float), ('nms_keypoint_threshold', C.c_float),
        ('nms_suppression', C.c_float), ('cif_min_scale', C.c_float),
        ('caf_min_distance', C.c_float), ('caf_max_distance', C.c_float),
        ('connection_method', C.c_int32), ('greedy', C.c_int32),
        ('force_complete', C.c_int32), ('nms', C.c_int32), ('exp_mode', C.c_int32),
        ('seed_mask', C.c_uint32), ('profile', C.c_void_p), ('initial', C.POINTER(Initial)),
        ('inverse', C.POINTER(Inverse)), ('host_mapped', C.c_int32), ('in_flight', C.c_int32),
        ('asm_heavy_seeds', C.c_int32), ('asm_stage_kb', C.c_int32), ('cifhr_group', C.c_int32),
        ('cifhr_ctas_per_sm', C.c_int32), ('cifhr_rshift', C.c_int32), ('cif_neighbors', C.c_float),
        ('seed_mask_enabled', C.c_int32), ('cifhr_warps', C.c_int32), ('cifhr_depth', C.c_int32), ('splat_gate', C.c_void_p),
    ]


class Skeleton(C.Structure):
    _fields_ = [
        ('L', C.c_int32),
        ('joints', (C.c_int32 * 2) * PP_MAX_L),
        ('has_confidence_scales', C.c_int32),
        ('confidence_scales', C.c_float * PP_MAX_L),
    ]

    @classmethod
    def from_list(cls, skeleton, confidence_scales=None):
        if len(skeleton) > PP_MAX_L:
            raise ValueError('skeleton has %d links, library limit is %d' % (len(skeleton), PP_MAX_L))
        sk = cls()
        sk.L = len(skeleton)
        for i, (a, b) in enumerate(skeleton):
            sk.joints[i][0], sk.joints[i][1] = int(a), int(b)
        if confidence_scales is not None:
            sk.has_confidence_scales = 1
            for i, v in enumerate(confidence_scales):
                sk.confidence_scales[i] = float(v)
        return sk


_OUT_FIELDS = ('hr', 'seeds', 'n_seeds', 'caf_fwd', 'caf_bwd', 'n_caf_fwd', 'n_caf_bwd',
               'ann_data', 'ann_scales', 'ann_score', 'ann_source', 'n_ann', 'pre_data',
               'pre_scales', 'pre_order', 'n_pre', 'status', 'packed')


class Outputs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _OUT_FIELDS]


class PifPafError(RuntimeError):
    def __init__(self, code, detail=''):
        self.code = code
        super().__init__('libpifpaf_b200 error %d: %s%s' % (code, error_string(code),
                                                            (' -- ' + detail) if detail else ''))


_lib = None


def lib():
    """Load the CUDA library (once). Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            'libpifpaf_b200.so is not built (%s). Run `python -c "import __graft_entry__ as g; '
            'g.build()"` or `python openpifpaf_b200/build.py`. There is no CPU fallback.' % LIB_PATH)
    l = C.CDLL(LIB_PATH)
    vp, sz, i32, f32 = C.c_void_p, C.c_size_t, C.c_int, C.c_float
    P = C.POINTER
    l.pp_error_string.restype = C.c_char_p
    l.pp_error_string.argtypes = [i32]
    l.pp_last_cuda_error.restype = C.c_char_p
    l.pp_version.restype = C.c_char_p
    l.pp_profile_create.restype = vp
    l.pp_profile_create.argtypes = [i32]
    l.pp_profile_destroy.restype = None
    l.pp_profile_destroy.argtypes = [vp]
    l.pp_profile_reset.restype = None
    l.pp_profile_reset.argtypes = [vp]
    l.pp_profile_read.argtypes = [vp, vp, vp]
    l.pp_profile_read_abs.argtypes = [vp, vp, vp]
    l.pp_stage_name.restype = C.c_char_p
    l.pp_stage_name.argtypes = [i32]
    l.pp_decode_launch_count.argtypes = [P(Params)]
    l.pp_default_params.restype = None
    l.pp_default_params.argtypes = [P(Params)]
    l.pp_default_caps.argtypes = [P(Shape), P(Caps)]
    l.pp_coco_skeleton.argtypes = [P(Skeleton)]
    l.pp_workspace_bytes.restype = sz
    l.pp_workspace_bytes.argtypes = [P(Shape), P(Caps)]
    l.pp_det_default_params.argtypes = [P(DetParams)]
    l.pp_det_default_params.restype = None
    l.pp_cifdet_workspace_bytes.argtypes = [P(Shape), P(Caps)]
    l.pp_cifdet_workspace_bytes.restype = sz
    l.pp_cifdet_decode_batch.argtypes = [vp, P(DetOutputs), vp, sz, P(Shape), P(Caps), P(DetParams), vp]
    l.pp_detections_inverse.argtypes = [vp, vp, P(Inverse), i32, i32, vp]
    l.pp_detection_nms_workspace_bytes.argtypes = [i32, i32]
    l.pp_detection_nms_workspace_bytes.restype = sz
    l.pp_detection_nms.argtypes = [vp] * 10 + [sz, i32, i32, P(DetParams), vp]
    l.pp_decode_batch_multiscale.argtypes = [P(FieldSet), i32, i32, P(Skeleton), P(Outputs), vp, sz,
                                             P(Shape), P(Caps), P(Params), vp]
    l.pp_sizeof.argtypes = [i32]
    l.pp_workspace_bytes_ex.restype = C.c_size_t
    l.pp_workspace_bytes_ex.argtypes = [C.POINTER(Shape), C.POINTER(Caps), i32]
    l.pp_hr_alloc.restype = C.c_int
    l.pp_hr_alloc.argtypes = [C.c_size_t, C.POINTER(vp), C.POINTER(i32)]
    l.pp_hr_free.restype = C.c_int
    l.pp_hr_free.argtypes = [vp]
    l.pp_gate_create.restype = vp
    l.pp_gate_create.argtypes = []
    l.pp_gate_destroy.restype = None
    l.pp_gate_destroy.argtypes = [vp]
    l.pp_packed_bytes.restype = sz
    l.pp_packed_bytes.argtypes = [P(Shape), P(Caps)]
    l.pp_packed_record_words.argtypes = [i32]
    l.pp_packed_header_words.restype = sz
    l.pp_packed_header_words.argtypes = [i32]
    l.pp_pack_results.argtypes = [P(Outputs), P(Shape), P(Caps), vp]
    l.pp_sizeof.restype = i32
    l.pp_annotations_inverse.argtypes = [vp, vp, vp, vp, vp, C.POINTER(Inverse), i32, i32, i32, vp]
    l.pp_head_epilogue.argtypes = [vp, vp, i32, i32, i32, i32, i32, i32, vp]
    l.pp_cifhr_accumulate.argtypes = [vp, vp, vp, sz, P(Shape), P(Caps), P(Params), vp, vp]
    l.pp_cif_seeds.argtypes = [vp, vp, vp, vp, vp, sz, P(Shape), P(Caps), P(Params), vp, vp]
    l.pp_caf_scored.argtypes = [vp, vp, P(Skeleton), f32, vp, vp, vp, vp, P(Shape), P(Caps),
                                P(Params), vp, vp]
    l.pp_cifcaf_assemble.argtypes = [vp] * 10 + [P(Skeleton), P(Outputs), vp, sz, P(Shape), P(Caps),
                                                 P(Params), vp]
    l.pp_keypoint_nms.argtypes = [vp, vp, vp, vp, vp, vp, sz, P(Shape), P(Caps), P(Params), vp]
    l.pp_cifhr_accumulate_ex.argtypes = [vp, vp, vp, i32, i32, vp, sz, P(Shape), P(Shape), P(Caps), P(Params),
                                         vp, vp]
    l.pp_keypoint_nms_noalloc.argtypes = [vp, vp, vp, vp, vp, P(Shape), P(Caps), P(Params), vp]
    l.pp_keypoint_nms_noalloc_max_bytes.restype = sz
    l.pp_keypoint_nms_noalloc_max_bytes.argtypes = []
    l.pp_keypoint_nms_noalloc_bytes.restype = sz
    l.pp_keypoint_nms_noalloc_bytes.argtypes = [P(Shape), P(Caps)]
    l.pp_decode_batch.argtypes = [vp, vp, P(Skeleton), P(Outputs), vp, sz, P(Shape), P(Caps),
                                  P(Params), vp]
    l.pp_decode_batch_host.argtypes = [vp, vp, vp, vp, P(Skeleton), P(Outputs), P(Outputs), vp, sz,
                                       P(Shape), P(Caps), P(Params), vp]
    l.pp_pipeline_create.restype = vp
    l.pp_pipeline_create.argtypes = [i32]
    l.pp_pipeline_destroy.restype = None
    l.pp_pipeline_destroy.argtypes = [vp]
    l.pp_pipeline_wait.argtypes = [vp, i32]
    l.pp_pipeline_chunks.argtypes = [vp]
    l.pp_decode_batch_host_async.argtypes = [vp, vp, vp, vp, vp, P(Skeleton), P(Outputs), P(Outputs),
                                             vp, sz, P(Shape), P(Caps), P(Params), i32, vp]
    l.pp_scalar_square_add_gauss_with_max.argtypes = [vp, i32, i32, vp, vp, vp, vp, i32, f32, f32, vp]
    l.pp_scalar_values.argtypes = [vp, i32, i32, vp, vp, i32, f32, vp, vp]
    l.pp_caf_center_s.argtypes = [vp, i32, f32, f32, f32, vp, vp, vp]
    l.pp_scalar_nonzero_clipped_with_reduction.argtypes = [vp, i32, i32, f32, f32, f32, vp, vp]
    l.pp_scalar_square_add_single.argtypes = [vp, i32, i32, C.c_longlong, C.c_longlong,
                                              C.c_longlong, i32, vp]
    for which, mirror in enumerate((Params, ImageMeta, Outputs, Skeleton, Shape, Caps, DetParams,
                                    DetOutputs)):
        if l.pp_sizeof(which) != C.sizeof(mirror):
            raise ImportError('ctypes mirror %s (%d bytes) does not match the library (%d): rebuild' % (
                mirror.__name__, C.sizeof(mirror), l.pp_sizeof(which)))
    _lib = l
    return l


EXPORTS = (
    'pp_profile_create', 'pp_profile_destroy', 'pp_profile_reset', 'pp_profile_read',
    'pp_stage_name', 'pp_decode_launch_count', 'pp_profile_read_abs', 'pp_gate_create', 'pp_gate_destroy',
    'pp_error_string', 'pp_last_cuda_error', 'pp_version', 'pp_default_params', 'pp_default_caps',
    'pp_coco_skeleton', 'pp_workspace_bytes', 'pp_workspace_bytes_ex', 'pp_hr_alloc', 'pp_hr_free', 'pp_sizeof', 'pp_decode_batch_multiscale', 'pp_det_default_params', 'pp_cifdet_workspace_bytes', 'pp_cifdet_decode_batch', 'pp_detections_inverse', 'pp_detection_nms_workspace_bytes', 'pp_detection_nms',
    'pp_head_epilogue', 'pp_annotations_inverse', 'pp_cifhr_accumulate', 'pp_cif_seeds',
    'pp_caf_scored', 'pp_cifcaf_assemble', 'pp_keypoint_nms', 'pp_decode_batch',
    'pp_keypoint_nms_noalloc', 'pp_keypoint_nms_noalloc_max_bytes', 'pp_keypoint_nms_noalloc_bytes',
    'pp_cifhr_accumulate_ex', 'pp_packed_bytes', 'pp_packed_record_words', 'pp_packed_header_words', 'pp_pack_results',
    'pp_decode_batch_host', 'pp_decode_batch_host_async', 'pp_pipeline_create',
    'pp_pipeline_destroy', 'pp_pipeline_wait', 'pp_pipeline_chunks', 'pp_scalar_square_add_gauss_with_max', 'pp_scalar_values',
    'pp_caf_center_s', 'pp_scalar_nonzero_clipped_with_reduction', 'pp_scalar_square_add_single',
)


def error_string(code):
    try:
        return lib().pp_error_string(code).decode()
    except Exception:  # pylint: disable=broad-except
        return 'code %d' % code


def check(code):
    if code != PP_OK:
        detail = ''
        if code == PP_ERR_CUDA:
            detail = lib().pp_last_cuda_error().decode()
        raise PifPafError(code, detail)

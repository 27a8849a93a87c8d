"""Command-line flags and decoder construction.

Same entry points, flag names, defaults and class-attribute side effects as the
reference's ``decoder/factory.py:17-213`` so that ``predict`` / ``eval_coco`` /
``video``-style callers work unchanged.  Two deliberate differences:

* ``factory_decode(dense_connections=True)`` builds the 44-link skeleton from a
  COPY.  The reference appends to ``head_nets[1].meta.skeleton`` in place and
  thereby mutates the global COCO_PERSON_SKELETON (decoder/factory.py:182-188).
* ``--decoder-workers`` is parsed for compatibility and ignored: batches are
  decoded on the GPU, there is no process pool.

``dense_coupling`` follows the reference exactly, including that it never
reaches ``CifCaf`` (it is stored in ``field_config.confidence_scales`` only).
"""
import logging

from . import nms
from .caf_scored import CafScored
from .cif_hr import CifHr
from .cif_seeds import CifSeeds
from .field_config import FieldConfig
from .generator.cifcaf import CifCaf
from .generator.cifdet import CifDet

LOG = logging.getLogger(__name__)


def cli(parser, *,
        force_complete_pose=True,
        seed_threshold=0.2,
        instance_threshold=0.0,
        keypoint_threshold=None,
        workers=None):
    group = parser.add_argument_group('decoder configuration')
    group.add_argument('--seed-threshold', default=seed_threshold, type=float,
                       help='minimum threshold for seeds')
    group.add_argument('--instance-threshold', type=float, default=instance_threshold,
                       help='filter instances by score')
    group.add_argument('--keypoint-threshold', type=float, default=keypoint_threshold,
                       help='filter keypoints by score')
    group.add_argument('--decoder-workers', default=workers, type=int,
                       help='accepted for compatibility; decoding is batched on the GPU')
    group.add_argument('--dense-connections', default=False, action='store_true',
                       help='use dense connections')
    group.add_argument('--dense-coupling', default=0.01, type=float, help='dense coupling')
    group.add_argument('--caf-seeds', default=False, action='store_true', help='[experimental]')
    if force_complete_pose:
        group.add_argument('--no-force-complete-pose', dest='force_complete_pose',
                           default=True, action='store_false')
    else:
        group.add_argument('--force-complete-pose', dest='force_complete_pose',
                           default=False, action='store_true')
    group.add_argument('--profile-decoder', nargs='?', const='profile_decoder.prof', default=None,
                       help='specify out .prof file or nothing for default file name')

    group = parser.add_argument_group('CifCaf decoders')
    group.add_argument('--cif-th', default=CifHr.v_threshold, type=float, help='cif threshold')
    group.add_argument('--caf-th', default=CafScored.default_score_th, type=float,
                       help='caf threshold')
    group.add_argument('--connection-method', default=CifCaf.connection_method,
                       choices=('max', 'blend'), help='connection method to use, max is faster')
    group.add_argument('--greedy', default=False, action='store_true', help='greedy decoding')


def configure(args):
    """Write the parsed flags into the class attributes the decoder reads at call time."""
    if args.keypoint_threshold is None:
        args.keypoint_threshold = 0.001 if not args.force_complete_pose else 0.0
    if args.force_complete_pose:
        assert args.keypoint_threshold == 0.0
    assert args.seed_threshold >= args.keypoint_threshold

    CifHr.v_threshold = args.cif_th
    CifSeeds.threshold = args.seed_threshold
    CafScored.default_score_th = args.caf_th

    CifCaf.force_complete = args.force_complete_pose
    CifCaf.keypoint_threshold = args.keypoint_threshold
    CifCaf.greedy = args.greedy
    CifCaf.connection_method = args.connection_method

    nms.Detection.instance_threshold = args.instance_threshold
    nms.Keypoints.instance_threshold = args.instance_threshold
    nms.Keypoints.keypoint_threshold = args.keypoint_threshold

    if args.decoder_workers is None and getattr(args, 'batch_size', 1) > 1 \
       and not getattr(args, 'debug', False):
        args.decoder_workers = args.batch_size


def factory_from_args(args, model):
    configure(args)
    decode = factory_decode(model.head_nets,
                            basenet_stride=model.base_net.stride,
                            dense_coupling=args.dense_coupling,
                            dense_connections=args.dense_connections,
                            caf_seeds=args.caf_seeds,
                            multi_scale=getattr(args, 'multi_scale', False),
                            multi_scale_hflip=getattr(args, 'multi_scale_hflip', True),
                            worker_pool=args.decoder_workers)
    if getattr(args, 'profile_decoder', None) is not None:
        from .profiler import Profiler
        decode.__class__.__call__ = Profiler(decode.__call__, out_name=args.profile_decoder)
    return decode


def _is_intensity(meta):
    return hasattr(meta, 'keypoints') and not hasattr(meta, 'skeleton')


def _is_association(meta):
    return hasattr(meta, 'keypoints') and hasattr(meta, 'skeleton')


def factory_decode(head_nets, *,
                   basenet_stride,
                   dense_coupling=0.0,
                   dense_connections=False,
                   caf_seeds=False,
                   multi_scale=False,
                   multi_scale_hflip=True,
                   worker_pool=None):
    """Instantiate a decoder for the given head networks."""
    assert not caf_seeds, 'not implemented'
    head_names = tuple(hn.meta.name for hn in head_nets)
    LOG.debug('head names = %s', head_names)

    if hasattr(head_nets[0].meta, 'categories'):          # DetectionMeta, factory.py:138-149
        field_config = FieldConfig()       # the reference keeps the default stride 8 here
        return CifDet(field_config, head_nets[0].meta.categories, worker_pool=worker_pool)

    if len(head_nets) >= 2 and _is_intensity(head_nets[0].meta) \
       and _is_association(head_nets[1].meta):
        field_config = FieldConfig()
        if multi_scale:
            # factory.py:153-180, kept as the reference has it: head sets of 3 (cif, caf, caf25), or
            # of 2 with dense connections, five scales, doubled with the hflip heads.  Like the
            # reference this indexes head_nets by the multi-scale head index, so it needs a
            # head_nets list that long (IndexError otherwise); a FieldConfig built by hand with the
            # true per-scale strides is the working route (CifCaf.decode_batch_multiscale).
            per_set = 2 if dense_connections else 3
            n_scales = 10 if multi_scale_hflip else 5
            field_config.cif_indices = [v * per_set for v in range(n_scales)]
            field_config.caf_indices = [v * per_set + 1 for v in range(n_scales)]
            field_config.cif_min_scales = [0.0, 12.0, 16.0, 24.0, 40.0] * (n_scales // 5)
            field_config.caf_min_distances = [v * 3.0 for v in field_config.cif_min_scales]
            field_config.caf_max_distances = [160.0, 240.0, 320.0, 480.0, None] * (n_scales // 5)
        field_config.cif_strides = [head_nets[i].stride(basenet_stride)
                                    for i in field_config.cif_indices]
        field_config.caf_strides = [head_nets[i].stride(basenet_stride)
                                    for i in field_config.caf_indices]
        out_skeleton = head_nets[1].meta.skeleton
        skeleton = list(out_skeleton)
        if dense_connections:
            extra = list(head_nets[2].meta.skeleton)
            field_config.confidence_scales = (
                [1.0 for _ in skeleton] + [dense_coupling for _ in extra])
            skeleton = skeleton + extra
            # the reference's in-place append makes out_skeleton the long list too
            out_skeleton = skeleton
        return CifCaf(
            field_config,
            keypoints=head_nets[0].meta.keypoints,
            skeleton=skeleton,
            out_skeleton=out_skeleton,
            worker_pool=worker_pool,
        )

    raise Exception('decoder unknown for head names: {}'.format(head_names))

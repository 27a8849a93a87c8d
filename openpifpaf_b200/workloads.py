"""Workload tooling: fields of a randomly initialised resnet50 + CifCaf heads
(BASELINE.json configs[1], SURVEY.md 8d "config 2").  Not on the decode path.

The reference builds this network with ``network.factory(base_name='resnet50',
head_names=['cif', 'caf'], pretrained=False)``: torchvision's resnet50 without
its max-pool (``network/factory.py:266-300``, ``basenetworks.py:27-57``; output
stride 16, 2048 features) followed by one ``CompositeFieldFused`` per head
(``network/heads.py:363-404``: a 1x1 convolution to ``n_fields * (1 + 4
n_vectors) * 4`` channels, default initialisation).  The same architecture is
assembled here from torchvision and ``torch.nn.Conv2d`` so that the GPU arm of
``bench.py`` does not need the staged reference; the head epilogue (pixel
shuffle, sigmoid / exp, index add, CAF reorder) is this package's
``network.CifCafCollector`` (csrc/head.cu).
"""
import torch

N_KEYPOINTS, N_LINKS = 17, 19


def build_random_init_net(seed=0, n_keypoints=N_KEYPOINTS, n_links=N_LINKS):
    """(trunk, conv_cif, conv_caf) in eval mode, default initialisation under ``seed``."""
    import torchvision
    torch.manual_seed(seed)
    vision = torchvision.models.resnet50(weights=None)
    mods = list(vision.children())
    trunk = torch.nn.Sequential(mods[0], mods[1], mods[2], mods[4], mods[5], mods[6], mods[7])
    conv_cif = torch.nn.Conv2d(2048, n_keypoints * 5 * 4, 1)
    conv_caf = torch.nn.Conv2d(2048, n_links * 9 * 4, 1)
    return trunk.eval(), conv_cif.eval(), conv_caf.eval()


@torch.no_grad()
def random_init_head_outputs(batch, image_hw=(641, 641), *, seed=0, device='cuda', image_seed=None,
                             micro_batch=16):
    """Raw head convolution outputs ``(B, 340, h, w)``, ``(B, 684, h, w)`` for
    ``torch.randn`` images (641 px -> h = w = 41)."""
    trunk, conv_cif, conv_caf = (m.to(device) for m in build_random_init_net(seed))
    gen = torch.Generator(device='cpu').manual_seed(seed + 1 if image_seed is None else image_seed)
    xs_cif, xs_caf = [], []
    for b0 in range(0, batch, micro_batch):
        n = min(micro_batch, batch - b0)
        images = torch.randn((n, 3) + tuple(image_hw), generator=gen).to(device)
        feat = trunk(images)
        xs_cif.append(conv_cif(feat).float())
        xs_caf.append(conv_caf(feat).float())
    return torch.cat(xs_cif).contiguous(), torch.cat(xs_caf).contiguous()


def collect_torch(x, n_fields, n_vectors):
    """Plain torch statement of the eval-mode head epilogue + collector for one head
    (``network/heads.py:406-448,65-88``), used where no CUDA device runs the kernel
    (the reference arm's field generation on the host)."""
    B = x.shape[0]
    x = torch.nn.functional.pixel_shuffle(x, 2)[:, :, :-1, :-1]
    H, W = x.shape[2:]
    g = [n_fields, n_fields * n_vectors * 2, n_fields * n_vectors, n_fields * n_vectors]
    conf, reg, logb, scale = torch.split(x, g, dim=1)
    conf = torch.sigmoid(conf).view(B, n_fields, 1, H, W)
    reg = reg.reshape(B, n_fields, n_vectors, 2, H, W).clone()
    logb = logb.reshape(B, n_fields, n_vectors, H, W)
    scale = torch.exp(scale).view(B, n_fields, n_vectors, H, W)
    ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32, device=x.device),
                            torch.arange(W, dtype=torch.float32, device=x.device), indexing='ij')
    reg[:, :, :, 0] += xs
    reg[:, :, :, 1] += ys
    parts = [conf]
    for v in range(n_vectors):
        parts += [reg[:, :, v], logb[:, :, v:v + 1], scale[:, :, v:v + 1]]
    return torch.cat(parts, dim=2).contiguous()


def random_init_head_fields(batch, image_hw=(641, 641), *, seed=0, image_seed=None, device='cuda',
                            max_scale=None):
    """``(cif (B,17,5,H,W), caf (B,19,9,H,W))`` float32 on ``device``.

    ``max_scale``: the documented *tamed* variant of SURVEY 8d config 2 -- the scale
    channels are clamped to that many cells (the raw ones reach 1e18)."""
    x_cif, x_caf = random_init_head_outputs(batch, image_hw, seed=seed, device=device, image_seed=image_seed)
    if torch.device(device).type == 'cuda':
        from .network import FusedHeadEpilogue
        cif = FusedHeadEpilogue(N_KEYPOINTS, 1)(x_cif)
        caf = FusedHeadEpilogue(N_LINKS, 2)(x_caf)
    else:
        cif = collect_torch(x_cif, N_KEYPOINTS, 1)
        caf = collect_torch(x_caf, N_LINKS, 2)
    if max_scale is not None:
        cif[:, :, 4].clamp_(max=max_scale)
        caf[:, :, 4].clamp_(max=max_scale)
        caf[:, :, 8].clamp_(max=max_scale)
    return cif, caf


def field_statistics(cif, threshold=0.1, stride=8):
    """Active fraction, scale percentiles and splat-pixel updates per image of a CIF tensor."""
    conf, x, y, scale = cif[:, :, 0], cif[:, :, 1], cif[:, :, 2], cif[:, :, 4]
    Hh, Wh = (cif.shape[3] - 1) * stride + 1, (cif.shape[4] - 1) * stride + 1
    active = conf > threshold
    s = scale[active].double()
    sigma = torch.clamp(0.5 * scale * stride, min=1.0)
    x0 = torch.clamp(x * stride - sigma, 0, Wh)
    x1 = torch.clamp(x * stride + sigma + 1, 0, Wh)
    y0 = torch.clamp(y * stride - sigma, 0, Hh)
    y1 = torch.clamp(y * stride + sigma + 1, 0, Hh)
    box = ((x1 - x0).clamp(min=0) * (y1 - y0).clamp(min=0)).double()
    q = torch.quantile(s[:1 << 20], torch.tensor([0.01, 0.5, 0.99], dtype=torch.float64, device=s.device))
    return {
        'active_fraction': float(active.float().mean()),
        'scale_percentiles_1_50_99_max': [float(v) for v in q] + [float(s.max())],
        'splat_pixel_updates_per_image': float((box * active).sum() / cif.shape[0]),
    }

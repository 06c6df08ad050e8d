"""Random-init UNet2d / UNet3d (bioimage.io nucleus-segmentation architecture).

Plain torch: the network forward is the one part of the path that stays a library call (tensor cores live here,
not in the kernels).  Encoder of `depth` levels (two 3x3 conv + ReLU, max-pool 2), features f*2^i, bottleneck,
decoder (x2 upsample + 1x1 conv, concat skip, conv block), 1x1 head.  fp32, eval mode, fixed seed.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def _block(nd: int, cin: int, cout: int) -> nn.Sequential:
    conv = nn.Conv2d if nd == 2 else nn.Conv3d
    return nn.Sequential(conv(cin, cout, 3, padding=1), nn.ReLU(inplace=True), conv(cout, cout, 3, padding=1),
                         nn.ReLU(inplace=True))


class _UNet(nn.Module):
    nd = 2

    def __init__(self, in_channels: int = 1, out_channels: int = 1, depth: int = 4, initial_features: int = 16,
                 final_sigmoid: bool = False):
        super().__init__()
        nd = self.nd
        conv = nn.Conv2d if nd == 2 else nn.Conv3d
        self.depth = depth
        feats = [initial_features * 2**i for i in range(depth + 1)]
        self.encoders = nn.ModuleList([_block(nd, in_channels if i == 0 else feats[i - 1], feats[i]) for i in range(depth)])
        self.base = _block(nd, feats[depth - 1], feats[depth])
        self.reducers = nn.ModuleList([conv(feats[i + 1], feats[i], 1) for i in reversed(range(depth))])
        self.decoders = nn.ModuleList([_block(nd, 2 * feats[i], feats[i]) for i in reversed(range(depth))])
        self.head = conv(feats[0], out_channels, 1)
        self.final_sigmoid = final_sigmoid

    def features(self, x: torch.Tensor) -> torch.Tensor:
        """Everything up to (not including) the 1x1 head."""
        pool = F.max_pool2d if self.nd == 2 else F.max_pool3d
        mode = "bilinear" if self.nd == 2 else "trilinear"
        skips = []
        for enc in self.encoders:
            x = enc(x)
            skips.append(x)
            x = pool(x, 2)
        x = self.base(x)
        for red, dec in zip(self.reducers, self.decoders):
            x = red(F.interpolate(x, scale_factor=2, mode=mode, align_corners=False))
            x = dec(torch.cat([skips.pop(), x], dim=1))
        return x

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.head(self.features(x))
        return torch.sigmoid(x) if self.final_sigmoid else x


class UNet2d(_UNet):
    nd = 2


class UNet3d(_UNet):
    nd = 3


ARCHITECTURES = {"UNet2d": UNet2d, "UNet3d": UNet3d}


def calibrate_head(model: nn.Module, probe: torch.Tensor, foreground: float = 0.5) -> nn.Module:
    """Re-centre and re-scale the 1x1 head so that a fraction `foreground` of the logits of `probe` is positive (0.5: the
    median lands on 0) and their spread is of order one.

    A random-init UNet's output is its head bias plus a few hundredths of variation, so every pixel lands on the same
    side of the threshold and the mask is all-0 or all-1 — the stitch/label stages would have nothing to do.  This keeps
    the random weights and only applies an affine map to the output channel(s).
    """
    with torch.no_grad():
        y = model(probe.to(next(model.parameters()).device))
        med = torch.quantile(y.flatten()[:: max(1, y.numel() // (1 << 22))].float(), 1.0 - float(foreground))
        spread = (y - med).abs().mean().clamp_min(1e-12)
        model.head.weight.div_(spread)
        model.head.bias.sub_(med).div_(spread)
    return model


def build_model(name: str, seed: int = 0, **kwargs) -> nn.Module:
    """Random-init, seed-fixed, eval-mode model (there is no network for zoo weights)."""
    if name not in ARCHITECTURES:
        raise ValueError(f"unknown architecture {name!r}; known: {sorted(ARCHITECTURES)}")
    gen_state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    model = ARCHITECTURES[name](**kwargs).eval()
    torch.random.set_rng_state(gen_state)
    for p in model.parameters():
        p.requires_grad_(False)
    return model


def build_from_source(root, source: str, callable_name: str, **kwargs) -> nn.Module:
    """The architecture a model package ships as a python file (`architecture: {source: unet.py, callable: UNet}`; what
    bioimageio.core imports for `ArchitectureFromFile`): import the file from the package directory, call the class."""
    import importlib.util
    from pathlib import Path

    root = Path(root).resolve()
    path = (root / source).resolve()
    if root not in path.parents or not path.is_file():
        raise FileNotFoundError(f"architecture source {source!r} is not a file inside the model package {root}")
    spec = importlib.util.spec_from_file_location(f"_b200_model_{abs(hash(str(path)))}", path)
    module = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(module)
    if not hasattr(module, callable_name):
        raise ValueError(f"{source} does not define {callable_name!r}")
    model = getattr(module, callable_name)(**kwargs).eval()
    for p in model.parameters():
        p.requires_grad_(False)
    return model


def strict_fp32() -> None:
    """The path is specified in fp32: no TF32 inside the forward either."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False

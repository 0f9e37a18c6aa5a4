"""Checkpoint loading: sevenn/checkpoint.py:258-311 + scripts/backward_compatibility.py:13-176."""
import os
from typing import Dict, Tuple

import numpy as np
import torch

from . import irreps as ir
from .spec import ModelSpec, build_spec


def _ver(v) -> Tuple[int, ...]:
    out = []
    for tok in str(v).split('.')[:3]:
        digits = ''.join(ch for ch in tok if ch.isdigit())
        out.append(int(digits) if digits else 0)
    return tuple(out + [0] * (3 - len(out)))


def patch_old_config(cfg: dict) -> dict:
    """backward_compatibility.py:13-36: checkpoints <= 0.9 used un-normalised spherical harmonics."""
    cfg = dict(cfg)
    if _ver(cfg.get('version', '0.0.0')) < (0, 10, 0) and '_normalize_sph' not in cfg:
        cfg['_normalize_sph'] = False
    return cfg


def patch_state_dict(spec: ModelSpec, sd: Dict[str, torch.Tensor], version) -> Dict[str, torch.Tensor]:
    """backward_compatibility.py:74-176 (sort_old_convolution): checkpoints written before 0.11.0
    store the last radial-MLP layer in instruction *creation* order; permute its columns into the
    sorted-by-output order used now and flip the sign where the stored w3j is minus today's."""
    sd = dict(sd)
    if _ver(version) >= (0, 11, 0):
        return sd
    flipped = set()
    for L in spec.layers:
        conv = L.conv
        key = f'{conv.name}.weight_nn.layer{len(conv.hidden) - 2}.weight'
        ww = sd[key]
        new = torch.empty_like(ww)
        by_out = {p_idx: p for p_idx, p in enumerate(conv.paths)}
        # conv.paths is sorted by output index c == position in the sorted list
        for (o_old, mul, c, (l1, l2, l3)) in conv.paths_creation_order:
            p = by_out[c]
            blk = ww[:, o_old:o_old + mul]
            wkey = f'{conv.name}.convolution._compiled_main_left_right._w3j_{l1}_{l2}_{l3}'
            if l1 > 0 and l2 > 0 and l3 > 0 and wkey in sd:
                now = torch.tensor(ir.wigner_3j(l1, l2, l3), dtype=sd[wkey].dtype)
                if not torch.allclose(sd[wkey], now, atol=1e-6):
                    assert torch.allclose(sd[wkey], -now, atol=1e-6), wkey
                    blk = -blk
                    flipped.add(wkey)
            new[:, p.w_off:p.w_off + mul] = blk
        sd[key] = new
    for wkey in flipped:          # like the reference (`stct[conv_w3j_key] *= -1`): the stored buffers follow the weights
        sd[wkey] = -sd[wkey]
    return sd


def load_checkpoint(path: str):
    """Returns (ModelSpec, state_dict ready for Engine, config)."""
    if not os.path.isfile(path):
        raise FileNotFoundError(path)
    cp = torch.load(path, weights_only=False, map_location='cpu')
    cfg = patch_old_config(cp['config'])
    spec = build_spec(cfg)
    sd = patch_state_dict(spec, cp['model_state_dict'], cfg.get('version', '0.0.0'))
    missing = [k for k in spec.expected_keys() if k not in sd]
    assert not missing, f'checkpoint misses {missing}'
    # The returned pair is in today's layout: say so, so that handing (cfg, sd) back to the `(config, state_dict)`
    # entry points (SevenNetCalculator(file_type='model_instance'), deploy()) does not sort / flip a second time.
    if _ver(cfg.get('version', '0.0.0')) < (0, 11, 0):
        cfg['_checkpoint_version'] = cfg.get('version')
        cfg['version'] = '0.11.0'
    return spec, sd, cfg

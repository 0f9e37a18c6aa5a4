"""Checkpoint loading: sevenn/checkpoint.py:258-311 + scripts/backward_compatibility.py:13-176."""
import os
from typing import Dict, Tuple

import torch

from . import irreps as ir
from .spec import ModelSpec, build_spec


def _ver(v) -> Tuple[int, ...]:
    out = []
    for tok in str(v).split('.')[:3]:
        digits = ''.join(ch for ch in tok if ch.isdigit())
        out.append(int(digits) if digits else 0)
    return tuple(out + [0] * (3 - len(out)))


def _ver_str(version):
    """(first three components as a string, 4th component) -- backward_compatibility.py:158-166"""
    if not version:
        raise ValueError('No version found in config')
    vs = str(version).split('.')
    suffix = vs[3] if len(vs) == 4 else ''
    return '.'.join(vs[:3]), suffix


def needs_map(version) -> bool:
    """backward_compatibility.py:168-169.  The reference compares version STRINGS, so '0.9.x' is not < '0.10.0'
    (nor < '0.11.0' below): a 0.9.x checkpoint is neither renamed nor re-sorted by the reference, and a drop-in must do
    the same to give the same numbers.  No 0.9.x checkpoint is in the reference tree to pin this on."""
    return _ver_str(version)[0] < '0.10.0'


def needs_sort(version) -> bool:
    """backward_compatibility.py:174-175 (string comparison, like the reference; '0.11.0.dev0' is still sorted)."""
    ver, suffix = _ver_str(version)
    return ver < '0.11.0' or (ver == '0.11.0' and suffix == 'dev0')


def patch_old_config(cfg: dict, require_version: bool = True) -> dict:
    """backward_compatibility.py:13-36.  ``require_version=False``: a (config, state_dict) pair handed over in memory
    (not a checkpoint file) without a version is in today's layout."""
    cfg = dict(cfg)
    version = cfg.get('version', None)
    if not version:
        if not require_version:
            return cfg
        raise ValueError('No version found in config')
    major, minor = _ver(version)[:2]
    if major == 0 and minor <= 9:
        cf = cfg.get('cutoff_function')
        if isinstance(cf, dict) and cf.get('cutoff_function_name') == 'XPLOR':
            cfg['cutoff_function'] = {k: v for k, v in cf.items() if k != 'poly_cut_p_value'}
        if 'train_denominator' not in cfg:
            cfg['train_denominator'] = cfg.pop('train_avg_num_neigh', False)
        if cfg.pop('optimize_by_reduce', None) is False:
            raise ValueError('This checkpoint(optimize_by_reduce: False) is no longer supported')
        if 'conv_denominator' not in cfg:
            cfg['conv_denominator'] = 0.0
        if '_normalize_sph' not in cfg:
            cfg['_normalize_sph'] = False
    return cfg


def map_old_model(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """backward_compatibility.py:39-71: module names of checkpoints written before the 2024-04 renaming
    ('0 convolution' -> '0_convolution', 'denumerator' -> 'denominator', ...)."""
    names = {'EdgeEmbedding': 'edge_embedding', 'reducing nn input to hidden': 'reduce_input_to_hidden',
             'reducing nn hidden to energy': 'reduce_hidden_to_energy', 'rescale atomic energy': 'rescale_atomic_energy'}
    for i in range(10):
        for old in ('self connection intro', 'convolution', 'self interaction 2', 'equivariant gate'):
            names[f'{i} {old}'] = f'{i}_' + old.replace(' ', '_')
    out = {}
    for k, v in sd.items():
        head, _, follower = k.partition('.')
        follower = follower.replace('denumerator', 'denominator')
        out[(names[head] + '.' + follower) if head in names else k] = v
    return out


def patch_state_dict(spec: ModelSpec, sd: Dict[str, torch.Tensor], version) -> Dict[str, torch.Tensor]:
    """backward_compatibility.py:74-176 (sort_old_convolution): checkpoints written before 0.11.0
    store the last radial-MLP layer in instruction *creation* order; permute its columns into the
    sorted-by-output order used now and flip the sign where the stored w3j is minus today's."""
    sd = dict(sd)
    if needs_map(version):
        sd = map_old_model(sd)
    if not needs_sort(version):
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
    sd = patch_state_dict(spec, cp['model_state_dict'], cfg.get('version'))
    missing = [k for k in spec.expected_keys() if k not in sd]
    assert not missing, f'checkpoint misses {missing}'
    # The returned pair is in today's layout: say so, so that handing (cfg, sd) back to the `(config, state_dict)`
    # entry points (SevenNetCalculator(file_type='model_instance'), deploy()) does not sort / flip a second time.
    if needs_sort(cfg.get('version')):
        cfg['_checkpoint_version'] = cfg.get('version')
        cfg['version'] = '0.11.0'
    return spec, sd, cfg

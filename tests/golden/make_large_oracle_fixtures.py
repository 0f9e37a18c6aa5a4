"""TEST INFRASTRUCTURE.  fp64 oracle outputs at sizes where the oracle needs minutes and tens of GB, so the `-m gpu`
parity tests of the path the bench really times (tcgen05 3xTF32 GEMM arm, tabulated radial weights, >= 1,000 atoms)
can compare against them on the GPU box without re-running the oracle there.

    python tests/golden/make_large_oracle_fixtures.py            # writes tests/golden/large_oracle_*.npz

The arithmetic is oracle/sevenn_ref.py unchanged; only the convolution is wrapped in torch.utils.checkpoint so that the
per-edge messages ([E, mid] per path, kept for autograd) of one layer at a time are alive (41 GB -> 9 GB at 768 atoms).
Weights are NOT stored: the tests rebuild them from the same seed and compare the sha1 recorded here.
"""
import copy
import hashlib
import os
import sys
import time

import numpy as np
import torch
from torch.utils.checkpoint import checkpoint

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import sevenn_ref as S  # noqa: E402
from tests.structures import fluorite_hfo2, jittered_lips  # noqa: E402
from tests.test_gpu_parity import _random_model  # noqa: E402

CASES = {
    '7net0_hfo2_324': ('7net-0', lambda: fluorite_hfo2((3, 3, 3), a=4.8)),
    '7net0_hfo2_1500': ('7net-0', lambda: fluorite_hfo2((5, 5, 5), a=4.8)),
    'l3i5_lips_1000': ('7net-l3i5', lambda: jittered_lips(n_side=10)),
}


def state_sha1(sd):
    h = hashlib.sha1()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(np.ascontiguousarray(sd[k].detach().cpu().numpy()).tobytes())
    return h.hexdigest()


def _conv_forward_checkpointed(self, data):
    # same statements as oracle IrrepsConvolution.forward, inside one checkpointed function
    idx = data[self.key_edge_idx]

    def f(emb, x, filt):
        weight = self.weight_nn(emb)
        message = self.convolution(x[idx[1]], filt, weight)
        out = torch.zeros(len(x), message.shape[1], dtype=x.dtype)
        out.index_add_(0, idx[0], message)
        return out.div(self.denominator)

    data[S.X] = checkpoint(f, data[self.key_weight_input], data[S.X], data[self.key_filter], use_reentrant=False)
    return data


def main(which):
    S.IrrepsConvolution.forward = _conv_forward_checkpointed
    for tag in which:
        name, make = CASES[tag]
        cfg, model = _random_model(name)
        numbers, pos, cell = make()
        t0 = time.time()
        torch.set_default_dtype(torch.float64)
        g = S.atoms_to_graph(numbers, pos, cell, [True] * 3, cfg['cutoff'], dtype=torch.float64)
        ref = copy.deepcopy(model).double()(dict(g))
        torch.set_default_dtype(torch.float32)
        out = os.path.join(ROOT, 'tests', 'golden', f'large_oracle_{tag}.npz')
        np.savez_compressed(
            out, numbers=numbers, pos=pos, cell=cell, n_edges=g[S.EDGE_IDX].shape[1],
            energy=ref[S.E_TOT].item(), forces=ref[S.FORCE].detach().numpy(),
            stress=ref[S.STRESS].detach().numpy().reshape(6), atomic_energy=ref[S.AE].detach().numpy().ravel(),
            state_sha1=state_sha1(model.state_dict()), preset=name)
        print(tag, len(numbers), 'atoms', g[S.EDGE_IDX].shape[1], 'edges', f'{time.time() - t0:.0f} s',
              'E', ref[S.E_TOT].item(), 'max|F|', ref[S.FORCE].abs().max().item(), flush=True)


if __name__ == '__main__':
    main(sys.argv[1:] or list(CASES))

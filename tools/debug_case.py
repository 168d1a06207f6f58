import sys, numpy as np, torch
sys.path.insert(0, '.')
from tests.golden_inputs import PHYS_CASES, dense_from_spec, phys_inputs
from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
g = dict(np.load('tests/golden/reference_vectors.npz'))
i = int(sys.argv[1]) if len(sys.argv) > 1 else 4
case = PHYS_CASES[i]
pred, obs, rho = phys_inputs(case)
dev = 'cuda'
val, mp = phys_cons_raw(pred.to(dev), obs.to(dev), dense_from_spec(case['psf']).to(dev), rho.to(dev) if torch.is_tensor(rho) else rho,
    reduction=case['reduction'], padding=case['padding'], crop=case['crop'], robust=case['robust'], eps=case['eps'], return_map=True)
d = np.abs(mp.cpu().numpy() - g[f'phys{i}_map'])
print('val', val.cpu().numpy(), g[f'phys{i}_val'])
print('max map diff', d.max(), 'count>1e-5', (d > 1e-5).sum())
idx = np.argwhere(d > 1e-5)
print(idx[:40])

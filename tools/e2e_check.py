"""Development check (GPU box): CUDA library vs CPU oracle on full fast_binless / difference / eval_cv runs."""
import sys, time, numpy as np
sys.path.insert(0, '.')
from binless_b200 import cuda_lib
from oracle.loader import oracle_lib
from binless_b200.synth import synth

gpu, ora = cuda_lib(), oracle_lib()

def relerr(a, b):
    return float(np.max(np.abs(a - b) / (1.0 + np.abs(b)))) if a.size else 0.0

def run(n, D, seed, alpha, lam2, lam1, **kw):
    mat = synth(n, D, 5000, seed)
    t = time.time(); hg = gpu.create(mat, n); og = hg.run_signal(alpha, lam2, lam1, **kw); tg = time.time() - t
    t = time.time(); ho = ora.create(mat, n); oo = ho.run_signal(alpha, lam2, lam1, **kw); to = time.time() - t
    dg, do = og['diagnostics'], oo['diagnostics']
    print(f'n={n} D={D} kw={kw}: steps gpu={dg["nsteps"]} cpu={do["nsteps"]} iters gpu={list(dg["solve_iters"])} cpu={list(do["solve_iters"])} '
          f't_gpu={tg:.3f}s (dev {dg["ms_total"]:.1f} ms, solver {dg["ms_solver"]:.1f} ms) t_cpu={to:.3f}s launches={hg.kernel_launches()}', flush=True)
    for k in ('log_signal', 'phihat', 'weight', 'signal', 'binless', 'background', 'biasmat', 'decaymat'):
        print(f'   {k:10s} relerr {relerr(og[k], oo[k]):.3e}')
    print('   prec_abs', np.abs(dg['prec_abs'] - do['prec_abs']).max(), 'patch equal', np.array_equal(og['patchno'], oo['patchno']),
          'npatch', og['patchno'].max() + 1, oo['patchno'].max() + 1)
    print('   biases', np.abs(og['biases']['estimate'] - oo['biases']['estimate']).max(), 'decay beta', np.abs(og['decay']['beta'] - oo['decay']['beta']).max(),
          'lambda', og['decay']['lambda_'], oo['decay']['lambda_'], 'expo', np.abs(og['exposures']['estimate'] - oo['exposures']['estimate']).max())
    if D > 1:
        dgf = hg.run_difference(1, alpha, 3.38, 0.14); dof = ho.run_difference(1, alpha, 3.38, 0.14)
        print('   difference: iters', list(dgf['diagnostics']['solve_iters']), list(dof['diagnostics']['solve_iters']),
              ' '.join(f'{k} {relerr(dgf[k], dof[k]):.2e}' for k in ('difference', 'deltahat', 'weight', 'phi_ref')),
              'patch equal', np.array_equal(dgf['patchno'], dof['patchno']))
    else:
        cg = hg.eval_cv(alpha, [1.0, 5.0], 0.05, group=1); co = ho.eval_cv(alpha, [1.0, 5.0], 0.05, group=1)
        for a, b in zip(cg, co):
            print('   eval_cv:', ' '.join(f'{k} {relerr(a[k], b[k]):.2e}' for k in ('signal', 'binless', 'phihat', 'weight', 'log_signal')),
                  'patch equal', np.array_equal(a['patchno'], b['patchno']))
    hg.close(); ho.close()

run(2, 1, 3, 0.49, 2.5, 0.0)
run(5, 2, 3, 0.49, 2.5, 0.0)
run(33, 2, 3, 0.49, [2.53, 4.51], [0.13, 0.10])
run(200, 2, 1, 0.49, [2.53, 4.51], [0.13, 0.10])
run(200, 1, 1, 0.49, 2.53, 0.1, tol_val=0.05, nouter=8)
run(131, 3, 5, 0.6, 1.5, 0.0, maxdiag=40)
run(257, 2, 6, 0.5, 2.5, 0.0, bg_steps=2, nouter=6, tol_val=0.01)
if len(sys.argv) > 1:
    n = int(sys.argv[1]); D = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    run(n, D, 2, 0.5, 2.5, 0.0)

import sys, time, numpy as np
sys.path.insert(0, '.')
from binless_b200._abi import FastBinlessLib, gfl_solve
from binless_b200.synth import synth
port = FastBinlessLib('oracle/_build/libfbo_port.so', 'fbo_')
gpu = FastBinlessLib('binless_b200/libfastbinless_cuda.so', 'fb_')
rng = np.random.default_rng(0)
def check(n, y, w, lam, conv, maxsteps=30000, maxdiag=0, label=''):
    t=time.time(); a = gfl_solve(port, n, y, w, lam, conv, maxsteps, maxdiag); tc=time.time()-t
    t=time.time(); b = gfl_solve(gpu, n, y, w, lam, conv, maxsteps, maxdiag); tg=time.time()-t
    print(f'{label} n={n} lam={lam} conv={conv} md={maxdiag}: steps cpu={a["steps"]} gpu={b["steps"]} alpha {a["alpha"]} {b["alpha"]} '
          f'|dbeta|={np.abs(a["beta"]-b["beta"]).max():.3e} |dz|={np.abs(a["z"]-b["z"]).max():.3e} |du|={np.abs(a["u"]-b["u"]).max():.3e} '
          f'|dfull|={np.abs(a["beta_full"]-b["beta_full"]).max():.3e} t_cpu={tc:.3f}s t_gpu={tg:.3f}s', flush=True)
    return a, b
for n in (2, 3, 5, 17, 33, 64, 100):
    nc = n*(n+1)//2
    y = rng.normal(0, 1, nc); w = rng.uniform(0, 3, nc); w[rng.random(nc) < 0.2] = 0
    check(n, y, w, 1.3, 1e-3, label='rand')
    check(n, y, w, 0.0, 1e-3, maxsteps=5, label='lam0')
# fixed iteration count (maxsteps) comparisons
n = 200; nc = n*(n+1)//2
y = rng.normal(0, 1, nc); w = rng.uniform(0, 3, nc)
for ms in (2, 3, 11, 52, 120):
    check(n, y, w, 2.5, 1e-12, maxsteps=ms, label='fixed-iters')
# realistic
mat = synth(200, 1, 5000, 1)
h = port.create(mat, 200)
out = h.run_signal(0.49, [2.53], 0.0)
a, b = check(200, out['phihat'], out['weight'], 2.53, 0.01, label='realistic')
a, b = check(200, out['phihat'], out['weight'], 2.53, 1e-4, label='realistic-tight')
a, b = check(200, out['phihat'], out['weight'], 2.53, 0.01, maxdiag=40, label='trapeze')
# warm start chain
st = gfl_solve(gpu, 200, out['phihat'], out['weight'], 2.53, 0.01)
st2 = gfl_solve(gpu, 200, out['phihat']*1.1, out['weight'], 4.0, 0.005, state=st)
sp = gfl_solve(port, 200, out['phihat'], out['weight'], 2.53, 0.01)
sp2 = gfl_solve(port, 200, out['phihat']*1.1, out['weight'], 4.0, 0.005, state=sp)
print('warm chain steps', st2['steps'], sp2['steps'], np.abs(st2['beta']-sp2['beta']).max())
if len(sys.argv) > 1:
    n = int(sys.argv[1]); nc = n*(n+1)//2
    mat = synth(n, 1, 5000, 2)
    h = port.create(mat, n)
    t=time.time(); out = h.run_signal(0.5, [2.5], 0.0); print('oracle full run', time.time()-t, out['diagnostics'])
    a, b = check(n, out['phihat'], out['weight'], 2.5, 0.01, label='realistic-big')
    for rep in range(3):
        t=time.time(); r = gfl_solve(gpu, n, out['phihat'], out['weight'], 2.5, 0.01); print('gpu solve wall', time.time()-t, r['steps'])

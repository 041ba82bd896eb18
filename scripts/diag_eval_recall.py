import sys, subprocess, os, numpy as np, tempfile, pathlib
sys.path.insert(0, 'tests')
from _oracle import write_vecbin, write_raw12
from _oracle import Oracle
oracle = Oracle()
tools = pathlib.Path('nano-vectordb_b200/bin')
tmp = pathlib.Path(tempfile.mkdtemp())
rng = np.random.default_rng(5)
n, d, k = 20_000, 64, 10
z = rng.standard_normal((n, 12)).astype(np.float32) @ rng.standard_normal((12, d)).astype(np.float32)
x = z + 0.2 * rng.standard_normal((n, d)).astype(np.float32)
x /= np.linalg.norm(x, axis=1, keepdims=True)
q = x[rng.integers(0, n, 200)] + 0.02 * rng.standard_normal((200, d)).astype(np.float32)
q /= np.linalg.norm(q, axis=1, keepdims=True)
write_vecbin(tmp / "base.vecbin", 2, oracle.to_f16(np.ascontiguousarray(x, np.float32)))
write_raw12(tmp / "q.raw12", np.ascontiguousarray(q, np.float32))
def run(*a, env=None):
    e = dict(os.environ); e.update(env or {})
    return subprocess.run([str(s) for s in a], capture_output=True, text=True, env=e)
env = {"IVF_NLIST": "64", "IVF_TRAIN": "10000", "PQ_M": "16", "IVF_NITER": "8"}
for b in range(2):
    r = run(tools / "nvdb_ivfpq_build", tmp / "base.vecbin", tmp / f"idx{b}.nvdbivf", env=env)
    print('build', r.returncode, __import__('hashlib').md5(open(tmp / f"idx{b}.nvdbivf",'rb').read()).hexdigest())
r = run(tools / "nvdb_gt_build", tmp / "base.vecbin", tmp / "q.raw12", k, tmp / "q.gtbin")
for rep in range(3):
    for e in ({"GT_PATH": str(tmp / "q.gtbin")}, {}, {"EXACT_METRIC": "DOT"}, {"PIPELINE": "interleaved"}):
        ee = {"IVF_NPROBE": "16", "REFINE_K": "50"}; ee.update(e)
        r = run(tools / "nvdb_ivf_eval", tmp / "base.vecbin", tmp / "idx0.nvdbivf", tmp / "q.raw12", k, env=ee)
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
        print(rep, list(e.keys()), [kv for kv in line[0].split() if kv.startswith('recall')] if line else r.stderr[-200:])

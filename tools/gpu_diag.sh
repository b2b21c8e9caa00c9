python - <<'PY'
import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import bench
import rpe_b200
from bench_support.synth import synth_batch, synth_K
B,W,H=128,1280,720
frames=synth_batch(3,0,B,W,H); K=synth_K(W,H)
host=torch.from_numpy(frames).pin_memory(); dev=host.cuda()
ecfg=rpe_b200.EngineConfig(max_width=W,max_height=H,max_pairs=B,nfeatures=5000,nlevels=12,match_mode="ratio",ratio=0.75,max_matches=500,ransac_max_iters=2000)
def run(n_eng, mode, depth=2, steps=20):
    engines=[rpe_b200.Engine(ecfg) for _ in range(n_eng)]
    streams=[torch.cuda.Stream() for _ in engines]
    for e,s in zip(engines,streams): e.set_stream(s.cuda_stream); e.set_pipeline(32,1)
    hn=host.numpy()
    sub = (lambda e: e.submit_device(dev.data_ptr(),H*W,B,H,W,W,K)) if mode=="dev" else (lambda e: e.submit_interleaved(hn,K))
    def steps_(n):
        q=[]
        for i in range(n):
            e=engines[i%len(engines)]
            if len(q)==len(engines)*depth: q.pop(0).wait()
            sub(e); q.append(e)
        for d in q: d.wait()
    steps_(4)
    res=[]
    for rep in range(3):
        torch.cuda.synchronize(); t0=time.perf_counter(); steps_(steps); torch.cuda.synchronize(); res.append((time.perf_counter()-t0)/steps*1e3)
    for e in engines: e.close()
    return [round(r,3) for r in res]
for n_eng, depth in ((1,1),(1,2),(2,1),(2,2)):
    for mode in ("dev","host"):
        print(n_eng, depth, mode, run(n_eng, mode, depth), flush=True)
PY

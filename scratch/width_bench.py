"""Row width of the code matrix: "auto" (64 / 128) against padding to 16, on C3-shaped fits with few features."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, b2rf
from b2rf import trainer
dev = torch.device("cuda:0")
n = 10_000_000
po = list(range(0, n + 1, 4096)) + ([n] if n % 4096 else [])
for F, strat in [(12, "all"), (24, "all"), (40, "onethird"), (100, "onethird")]:
    X, y, arity = b2rf.synth.c3(n, device=dev, F=F)
    for align in ["auto", 16]:
        trainer.CODE_ROW_ALIGN = align
        p = b2rf.TreeParams(maxDepth=10, maxBins=64, numTrees=50, bootstrap=True, seed=42, featureSubsetStrategy=strat)
        ts = []
        for rep in range(3):
            torch.cuda.synchronize(); t = time.perf_counter()
            f = trainer.fit_tensors(X, y, arity, p, part_offsets=po)
            torch.cuda.synchronize(); ts.append(time.perf_counter() - t)
        print(f"F={F} {strat} align={align} width={trainer.code_row_width(F)}: fit {min(ts[1:])*1e3:.1f} ms, K4 {f.timers['findBestSplits']*1e3:.1f} ms", flush=True)
    del X, y

import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch, numpy as np
from helpers import Golden, oracle_from_golden
from test_gpu_fused_search import build_search
from llm_guided_asr_b200 import synthetic as syn
g = Golden("normalize_length"); c = g.spec
dev = torch.device("cuda:0")
bs = build_search(g, dev)
xs = [torch.from_numpy(g["logp"]).to(dev)]
for seed, T in ((5, 50), (6, 72), (7, 33)):
    xs.append(syn.random_logp(seed, T, c["V"]).to(dev))
batch = bs.decode_batch(xs, c["maxr"], c["minr"])
orc = oracle_from_golden(g)
for b, x in enumerate(xs):
    alone = bs(x=x, maxlenratio=c["maxr"], minlenratio=c["minr"])
    ref = orc.forward(x.cpu().numpy(), c["maxr"], c["minr"])
    print("utt", b, "n:", len(batch[b]), len(alone), len(ref), "steps", bs.last_steps, orc.steps_run)
    for i, (hb, ha, hr) in enumerate(zip(batch[b], alone, ref)):
        ok = hb.yseq.tolist() == ha.yseq.tolist() == hr.yseq
        if not ok or i < 2:
            print(i, ok, float(hb.score), float(ha.score), float(hr.score), len(hb.yseq), len(ha.yseq), len(hr.yseq))
            if not ok:
                print("  batch", hb.yseq.tolist()); print("  alone", ha.yseq.tolist()); print("  orcl ", hr.yseq)
                break

"""Diagnostic: warm-start steps eager / graph x Adam-overlap on / off; prints which parameters differ from the reference run."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from bmhrl_b200 import ops, synth
from conftest import PAD, VOC, build_agent
from bmhrl_b200.model.bm_hrl_agent import BMManagerValueFunction, BMWorkerValueFunction
from bmhrl_b200.train import BMHRLTrainer, StepRunner

g = torch.load("tests/golden/reference_golden.pt", weights_only=False)
sd = synth.fill_state_dict(g["agent_shapes"], seed=1234)
b = synth.synthetic_batch(B=4, Sv=10, Sa=14, L=7, V=VOC, seed=2, ragged=True)
ops.set_math_mode(os.environ.get("BMHRL_MATH", "bf16"))
res = {}
CASES = (("eager/off", False, False, True, 255), ("graph/off", True, False, True, 255), ("graph/on", True, True, True, 255),
         ("graph/on#2", True, True, True, 255), ("graph/stage0", True, True, True, 1), ("graph/stage0#2", True, True, True, 1),
         ("graph/stage1", True, True, True, 2), ("graph/stage1#2", True, True, True, 2))
for name, use_graph, overlap, conc, stages in CASES:
    os.environ["BMHRL_EARLY_STAGES"] = str(stages)
    ops.CONCURRENT = conc
    ops.manual_seed(7)
    ops.RNG.salt = 5000
    model, cfg = build_agent(sd, dout_p=0.1)
    tr = BMHRLTrainer(model, BMWorkerValueFunction(cfg).cuda(), BMManagerValueFunction(cfg).cuda(), pad_idx=PAD)
    tr.overlap_grad_sync = overlap
    tr.teach_warmstart()
    model.manager.exploration = False
    run = StepRunner(tr, "warmstart", b, use_graph=use_graph, warmup=2 if use_graph else 0)
    n0 = 2 if use_graph else 0
    losses = [float(run.step(b)[0]) for _ in range(4 - n0)]
    torch.cuda.synchronize()
    res[name] = ({n: p.detach().clone() for n, p in model.named_parameters()}, losses, tr.cap_opt.step_counts.clone())
ref = res["eager/off"]
for name, (params, losses, steps) in res.items():
    bad = []
    for n in params:
        d = float((params[n] - ref[0][n]).norm() / (ref[0][n].norm() + 1e-30))
        if d > 1e-6:
            bad.append((n, d))
    print(name, "losses", [round(l, 5) for l in losses], "steps min/max", float(steps.min()), float(steps.max()), "differing:", len(bad))
    for n, d in sorted(bad, key=lambda t: -t[1])[:6]:
        print("    ", n, f"{d:.2e}")

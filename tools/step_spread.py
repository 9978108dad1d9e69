"""Run-to-run spread of the warm-start step's parameters after 3 steps: same configuration twice vs overlap on/off."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from bmhrl_b200 import ops, synth
from bmhrl_b200.model.bm_hrl_agent import BMManagerValueFunction, BMWorkerValueFunction
from bmhrl_b200.train import BMHRLTrainer, StepRunner
from conftest import PAD, VOC, build_agent, rel
from test_gpu_agent import _FakeSync

golden = torch.load(os.path.join(ROOT, "tests", "golden", "reference_golden.pt"), weights_only=False)
sd = synth.fill_state_dict(golden["agent_shapes"], seed=1234)
b = synth.synthetic_batch(B=4, Sv=10, Sa=14, L=7, V=VOC, seed=3, ragged=True)


def run(overlap, use_graph, conc=True):
    ops.CONCURRENT = conc
    ops.manual_seed(11, "cuda"); ops.RNG.salt = 1000
    model, cfg = build_agent(sd)
    tr = BMHRLTrainer(model, BMWorkerValueFunction(cfg).cuda(), BMManagerValueFunction(cfg).cuda(), pad_idx=PAD, grad_sync=_FakeSync())
    tr.overlap_grad_sync = overlap
    tr.teach_warmstart(); model.manager.exploration = False
    r = StepRunner(tr, "warmstart", b, use_graph=use_graph, warmup=1 if use_graph else 0)
    if use_graph: r._eager()
    else: r.step(b)
    for _ in range(2): r.step(b)
    torch.cuda.synchronize()
    ops.CONCURRENT = True
    return {k: v.detach().clone() for k, v in model.named_parameters() if v.requires_grad}


def worst(a, c):
    w = max(((rel(a[k], c[k]), k) for k in a), key=lambda t: t[0])
    q = "bm_enc.encoder.layers.0.self_att_M1.linear_Q2d.weight"
    return f"worst {w[0]:.2e} ({w[1]})  encQ {rel(a[q], c[q]):.2e}"


for ug in (True, False):
    base = run(False, ug)
    for i in range(3):
        print(f"graph={ug} same config (no overlap) rerun {i}: {worst(run(False, ug), base)}", flush=True)
    for i in range(3):
        print(f"graph={ug} overlap vs no overlap {i}:          {worst(run(True, ug), base)}", flush=True)
    ser = run(False, ug, conc=False)
    print(f"graph={ug} serial vs concurrent (no overlap):    {worst(ser, base)}", flush=True)
    print(f"graph={ug} serial rerun:                         {worst(run(False, ug, conc=False), ser)}", flush=True)

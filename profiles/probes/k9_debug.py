"""Stage isolation of K9 by weight surgery: which part of the pipeline disagrees with the float64 actor."""
import copy, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib
from distributional_rl_decision_and_control_b200.agent import Agent
from test_gpu_policy import _obs, _actor64
DEV = "cuda:0"
M = int(sys.argv[1]) if len(sys.argv) > 1 else 256
x = _obs(M, 1)

def run(name, edit):
    ag = Agent(device=DEV, seed=100, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
    a = ag.policy_local.actor
    with torch.no_grad():
        edit(a)
    want = _actor64(a)(tuple(t.double() for t in x))
    got = ag.act_batch(x, eps=0.0)
    torch.cuda.synchronize()
    try:
        _lib.policy_forward_status(); st = 0
    except Exception as e:
        st = str(e)
    err = (got.double() - want).abs()
    print(f"{name:34s} err {err.max().item():.3e}  rows_bad {(err.max(1).values > 1e-5).sum().item()}/{M} status {st}")
    bad = (err.max(1).values > 1e-5).nonzero().flatten()[:6].tolist()
    for r in bad[:3]:
        print("   row", r, "got", got[r].tolist(), "want", want[r].tolist())
    return got, want

z = lambda t: t.zero_()
run("all zero W (biases only)", lambda a: (z(a.hidden_layer.weight), z(a.hidden_layer_2.weight)))
run("W1=0, W2 random", lambda a: z(a.hidden_layer.weight))
run("W2=0, W1 random", lambda a: z(a.hidden_layer_2.weight))
def const_feats(a):
    z(a.self_encoder[0].weight); z(a.object_encoder[0].weight)
    a.self_encoder[0].bias.fill_(0.5); a.object_encoder[0].bias.fill_(0.25)
run("const features, W1 W2 random", const_feats)
def const_feats_w2zero(a):
    const_feats(a); z(a.hidden_layer_2.weight)
run("const features, W2=0", const_feats_w2zero)
def w1_self_only(a):
    a.hidden_layer.weight[:, 56:].zero_()
run("W1 uses self features only", w1_self_only)
def w1_obj_only(a):
    a.hidden_layer.weight[:, :56].zero_()
run("W1 uses object features only", w1_obj_only)
for c in range(4):
    def only_chunk(a, c=c):
        w = a.hidden_layer.weight.clone(); a.hidden_layer.weight.zero_(); a.hidden_layer.weight[:, 64*c:64*c+64] = w[:, 64*c:64*c+64]
    run(f"W1 chunk {c} only", only_chunk)
run("full", lambda a: None)

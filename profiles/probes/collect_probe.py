import torch, sys, os
sys.path.insert(0, os.getcwd())
from distributional_rl_decision_and_control_b200.agent import Agent
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv
dev = torch.device("cuda:0")
E = 8192
env = BatchedMarineNavEnv(E, 8, 10, 8, device=dev, seed=1, fused=True, prefetch="workers")
env.reset()
ag = Agent(device=dev, seed=100, agent_type="AC-IQN", BATCH_SIZE=1024, num_tau=32, BUFFER_SIZE=1 << 20)
def t(fn, n=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
eps = torch.tensor(0.05, device=dev)
o_self, o_obj, o_mask = env.obs
flat = (o_self.view(E * 8, 7), o_obj.view(E * 8, 5, 5), o_mask.view(E * 8, 5))
acts = ag.act_batch(flat, eps).view(E, 8, 2)
for tf32 in (False, True):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    print("tf32", tf32, "act_batch us", t(lambda: ag.act_batch(flat, eps)))
    print("   actor fwd only", t(lambda: ag.policy_local.actor(flat)))
    print("   tower only", t(lambda: ag.policy_local.actor._features(flat)))
print("step_fused us", t(lambda: env.step_fused(acts)))
out = env.step_fused(acts)
print("push us", t(lambda: ag.memory.push_step((o_self, o_obj, o_mask), acts, False, out, term=env.term)))

import sys, torch
sys.path.insert(0, '.')
from handyrl_b200.generation import BatchGenerator
N=65536
def timed(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/iters
gen = BatchGenerator("Geister", N, seed=1, depth=512, max_episodes=1<<20)
env=gen.env
main_w, scalar_w = env.alloc_obs(waves=2)
am=torch.empty(2, N, env.A, device="cuda"); act=torch.empty(2,N,dtype=torch.int32,device="cuda")
print("full            %.3f ms" % timed(lambda: gen.rollout(32, waves=dict(main=main_w, scalar=scalar_w, action_mask=am), out_action=act)))
print("compact only    %.3f ms" % timed(lambda: gen.rollout(32)))
print("obs only        %.3f ms" % timed(lambda: gen.rollout(32, waves=dict(main=main_w, scalar=scalar_w))))
print("amask only      %.3f ms" % timed(lambda: gen.rollout(32, waves=dict(action_mask=am))))

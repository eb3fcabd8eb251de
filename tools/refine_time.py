import sys; sys.path[:0]=['lightweight-mast3r-slam_b200','tests','oracle','.']
import torch, bench, synth, mast3r_slam_backends as be
sets = bench.make_match_sets('cuda', 8); M = synth.MATCH_DEFAULTS
for s in sets:
    p,_ = be.iter_proj(s['rays'], s['pts'], s['p_init'], 10, 1e-8, 1e-6); s['p1']=p.long()
ref = [be.refine_matches(s['D11'], s['D21'], s['p1'], 3, 5, _variant=1)[0] for s in sets[:2]]
got = [be.refine_matches(s['D11'], s['D21'], s['p1'], 3, 5)[0] for s in sets[:2]]
print('bit-exact vs direct kernel:', all(torch.equal(a,b) for a,b in zip(ref,got)))
def run(n):
    for i in range(n):
        s = sets[i % 8]; be.refine_matches(s['D11'], s['D21'], s['p1'], 3, 5)
run(20); torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); run(200); b.record(); torch.cuda.synchronize()
print('refine us', round(a.elapsed_time(b) / 200 * 1e3, 2))

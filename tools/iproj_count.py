import sys; sys.path[:0]=['lightweight-mast3r-slam_b200','tests','oracle','.']
import torch, bench, synth, mast3r_slam_backends as be
sets = bench.make_match_sets('cuda', 2)
for s in sets:
    p, c = be.iter_proj(s['rays'], s['pts'], s['p_init'], 10, 1e-8, 1e-6)
    raw = c.view(torch.uint8).int()
    conv, ev, last = raw & 1, (raw >> 1) & 15, (raw >> 5)   # last = (last fetching iteration)+1, 0 = none
    print("converged frac", conv.float().mean().item())
    print("evals per point: mean %.2f" % ev.float().mean().item(), "hist", torch.bincount(ev.flatten(), minlength=12).tolist())
    print("last fetching iteration+1 hist", torch.bincount(last.flatten(), minlength=11).tolist())
    w = last.view(-1, 32)
    print("per-warp (32 consecutive px) max last+1: mean %.2f" % w.max(1).values.float().mean().item(),
          "hist", torch.bincount(w.max(1).values, minlength=11).tolist())
    evw = ev.view(-1, 32)
    # evaluations a warp executes = number of iterations in which any lane fetched; upper bound by union, estimate by max last
    h, wd = 384, 512
    t = last.view(h, wd)
    blk = t.view(h // 4, 4, wd // 8, 8).permute(0, 2, 1, 3).reshape(-1, 32)
    print("per 8x4 patch max last+1: mean %.2f" % blk.max(1).values.float().mean().item())
    nc = (conv.view(h, wd) == 0)
    print("non-converged: frac %.3f, their mean evals %.2f; converged mean evals %.2f" % (nc.float().mean().item(), ev.view(h,wd)[nc].float().mean().item(), ev.view(h,wd)[~nc].float().mean().item()))

python - <<'P'
import sys, os
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"): sys.path.insert(0, os.path.join(os.getcwd(), p))
sys.path.insert(0, os.getcwd())
import torch, bench, statistics
import mast3r_slam_backends as be
sets = bench.make_match_sets("cuda", 8)
for s in sets:
    p, _ = be.iter_proj(s["rays"], s["pts"], s["p_init"], 10, 1e-8, 1e-6); s["p1"] = p.long()
for nb in (1, 2, 3, 4, 5):
    D11 = torch.cat([s["D11"] for s in sets[:nb]]); D21 = torch.cat([s["D21"] for s in sets[:nb]]); p1 = torch.cat([s["p1"] for s in sets[:nb]])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for v in (0, 3):
        for _ in range(3): be.refine_matches(D11, D21, p1, 3, 5, _variant=v)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(30)]
        for a, b in ev:
            flush.zero_()
            a.record(); be.refine_matches(D11, D21, p1, 3, 5, _variant=v); b.record()
        torch.cuda.synchronize()
        print(f"b={nb} refine variant {v}: median {statistics.median(a.elapsed_time(b) for a, b in ev) * 1e3:.1f} us")
P

cat > /tmp/san.py <<'P'
import sys, os
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(os.getcwd(), p))
import torch, synth
import mast3r_slam_backends as be
for (h, w, b) in [(96, 128, 2), (77, 101, 1), (200, 328, 1)]:
    m = synth.make_match_inputs(h, w, b=b, seed=12, device="cuda")
    D11, D21 = m["D11"].half(), m["D21"].view(b, -1, 24).half()
    p1 = m["true_uv"].round().long().clone()
    p1[..., 0] = (p1[..., 0] + w // 3).clamp(1, w - 2)
    p1[..., 1] = p1[..., 1].clamp(1, h - 2)
    p1 = p1.contiguous()
    a = be.refine_matches(D11, D21, p1, 3, 5)[0]
    c = be.refine_matches(D11, D21, p1, 3, 5, _variant=2)[0]
    torch.cuda.synchronize()
    print(h, w, b, torch.equal(a, c))
P
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --kernel-regex kns=refine python /tmp/san.py 2>&1 | tail -12
done

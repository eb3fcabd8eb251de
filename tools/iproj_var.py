"""Build variants of libm3r_b200.so with different iter_proj launch shapes (CPU box), or time them
(GPU box).  usage: iproj_var.py build | iproj_var.py time"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "lightweight-mast3r-slam_b200")
OUT = os.path.join(PKG, "build", "ipvar")
VARS = {"t128_c1": (128, 1), "t128_c8": (128, 8), "t128_c9": (128, 9), "t128_c10": (128, 10), "t64_c1": (64, 1),
        "t64_c16": (64, 16), "t64_c18": (64, 18), "t64_c21": (64, 21), "t256_c4": (256, 4), "t32_c32": (32, 32)}
if sys.argv[1] == "build":
    os.makedirs(OUT, exist_ok=True)
    objs = [os.path.join(PKG, "build", "libm3r_b200.so.o", f + ".cu.o") for f in ("common", "tracking", "pointmap", "gn")]
    for name, (t, c) in VARS.items():
        o = os.path.join(OUT, name + ".o")
        r = subprocess.run(["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC",
                        f"-DM3R_IP_THREADS={t}", f"-DM3R_IP_MINCTAS={c}", "-Xptxas", "-v", "-c", "-o", o,
                        os.path.join(PKG, "csrc", "matching.cu")], capture_output=True, text=True)
        lines = r.stderr.splitlines()
        for i, l in enumerate(lines):
            if "iter_proj_kernelILi2" in l and "Compiling" in l:
                print(name, lines[i + 2].strip(), "|", lines[i + 1].strip()[:60] if "spill" not in lines[i+2] else "")
                print("   ", [x for x in lines[i:i+4] if "spill" in x or "Used" in x])
        subprocess.check_call(["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o",
                               os.path.join(OUT, name + ".so"), o] + objs)
else:
    for name in VARS:
        code = ("import sys; sys.path[:0]=['lightweight-mast3r-slam_b200','tests','oracle','.']\n"
                "import torch, bench, synth, mast3r_slam_backends as be\n"
                "sets = bench.make_match_sets('cuda', 8); M = synth.MATCH_DEFAULTS\n"
                "def run(n):\n"
                "    for i in range(n):\n"
                "        s = sets[i % 8]; be.iter_proj(s['rays'], s['pts'], s['p_init'], 10, 1e-8, 1e-6)\n"
                "run(20); torch.cuda.synchronize()\n"
                "a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)\n"
                "a.record(); run(200); b.record(); torch.cuda.synchronize()\n"
                f"print('{name}', round(a.elapsed_time(b) / 200 * 1e3, 2), 'us')\n")
        env = dict(os.environ, M3R_B200_LIB=os.path.join(OUT, name + ".so"))
        subprocess.run([sys.executable, "-c", code], env=env, cwd=ROOT)

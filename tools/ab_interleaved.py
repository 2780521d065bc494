"""Same-box interleaved A/B of bench arms: every round runs every arm once, in a rotated
order, so that box drift (clocks, power state) spreads evenly over the arms.  Writes mean +- sd per arm.
usage: python tools/ab_interleaved.py ROUNDS OUT.json [arm ...]   (default: every arm)"""
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R1 = os.path.join(ROOT, "_ab", "r1")                     # round-1 tree (a092743), built beside this one; not in history
HEAD_CMD = [sys.executable, "bench.py", "--no-extra", "--no-cpu-baseline", "--steps", "20", "--warmup", "3"]
ARMS = {
    "head": (ROOT, HEAD_CMD, {}),
    "head_chunk_acc_off": (ROOT, HEAD_CMD, {"MIO_CHUNK_ACC": "off"}),
    "head_slot_align_1": (ROOT, HEAD_CMD, {"MIO_EXP_SLOT_ALIGN": "1"}),          # r1's 251-row slots (not batch invariant)
    "head_skip_norms": (ROOT, HEAD_CMD, {"MIO_EXP_SKIP_NORMS": "1"}),            # WRONG results: floor for any norm fusion
    "head_kb4": (ROOT, HEAD_CMD, {"MIO_CHUNK_KB": "4"}),                         # chunked accumulation, 4 k-blocks per chunk
    "head_kb6": (ROOT, HEAD_CMD, {"MIO_CHUNK_KB": "6"}),
    "head_gn_precise": (ROOT, HEAD_CMD, {"MIO_GN_PRECISE": "1"}),               # libdevice expf + IEEE divide in gn_apply_silu
    "r1": (R1, [sys.executable, "bench.py", "--no-cpu-baseline", "--no-latency", "--steps", "20", "--warmup", "3"], {}),
}


def run(name):
    cwd, cmd, env = ARMS[name]
    e = dict(os.environ); e.update(env)
    t = time.time()
    r = subprocess.run(cmd, cwd=cwd, env=e, capture_output=True, text=True, timeout=600)
    lines = [x for x in r.stdout.splitlines() if x.startswith("{")]
    if r.returncode != 0 or not lines:
        return {"error": (r.stderr or r.stdout)[-400:], "wall_s": time.time() - t}
    d = json.loads(lines[-1])
    return {"ms_per_step": d["ms_per_step"], "value": d["value"], "e2e": (d.get("e2e") or {}).get("value"),
            "sm_mhz": (d.get("clocks") or {}).get("sm_mhz"), "reasons": (d.get("clocks") or {}).get("reasons"),
            "wall_s": round(time.time() - t, 1)}


def main():
    rounds, out = int(sys.argv[1]), sys.argv[2]
    names = [n for n in (sys.argv[3:] or ["head", "head_chunk_acc_off", "head_slot_align_1", "head_skip_norms", "r1"]) if n != "r1" or os.path.exists(os.path.join(R1, "bench.py"))]
    res = {n: [] for n in names}
    for r in range(rounds):
        order = names[r % len(names):] + names[:r % len(names)]
        for n in order:
            x = run(n)
            res[n].append(x)
            print(r, n, x, flush=True)
    summary = {}
    for n, xs in res.items():
        ms = [x["ms_per_step"] for x in xs if "ms_per_step" in x]
        e2e = [x["e2e"] for x in xs if x.get("e2e")]
        summary[n] = {"n": len(ms), "ms_mean": statistics.mean(ms) if ms else None, "ms_sd": statistics.stdev(ms) if len(ms) > 1 else None,
                      "ms_min": min(ms) if ms else None, "e2e_mean": statistics.mean(e2e) if e2e else None,
                      "sm_mhz_mean": statistics.mean([x["sm_mhz"] for x in xs if x.get("sm_mhz")]) if any(x.get("sm_mhz") for x in xs) else None}
    json.dump({"summary": summary, "runs": res}, open(out, "w"), indent=1)
    for n, s in summary.items():
        print(n, s)


if __name__ == "__main__":
    main()

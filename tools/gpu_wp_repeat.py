"""Runs the pipeline program several times on the whole-program dataset of bench.py, reports every record that differs from
the unmodified reference's VCF and, with SVIPER_B200_TRACE=1, which inputs / outputs of the device job differ between runs
(gpurun -- python tools/gpu_wp_repeat.py [runs]). Outputs in gpurun_out/wp_*."""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import e2e_dataset  # noqa: E402

runs = int(sys.argv[1]) if len(sys.argv) > 1 else 6
out = os.path.join(ROOT, "gpurun_out")
os.makedirs(out, exist_ok=True)
tmp = tempfile.mkdtemp(prefix="svp_wp_")
d = e2e_dataset.make_c2(tmp, n_records=100, seed=2)
base = ["-c", d["vcf"], "-l", d["long"], "-s", d["short"], "-r", d["ref"], "-k", "400"]
subprocess.check_call([os.path.join(ROOT, "oracle", "_ref", "sviper")] + base + ["-t", "16", "-o", os.path.join(out, "wp_ref")],
                      stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
ref = [x for x in open(os.path.join(out, "wp_ref.vcf")) if not x.startswith("#")]
variants = [("default", {}), ("msahost", {"SVIPER_B200_MSA": "host"}), ("lanes1", {"SVP_LANES": "1"}),
            ("target64", {"SVP_BATCH_TARGET": "64", "SVP_LINGER_MS": "50"})]
traces = {}
for label, extra in variants:
    for r in range(runs if label == "default" else max(2, runs // 2)):
        env = dict(os.environ, SVIPER_B200_STATS="1", SVIPER_B200_TRACE="1")
        env.update(extra)
        name = "wp_%s_%d" % (label, r)
        try:
            p = subprocess.run([os.path.join(ROOT, "sviper_b200", "host", "sviper_b200")] + base + ["-t", "16", "-o", os.path.join(out, name)],
                               stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, env=env, timeout=90)
        except subprocess.TimeoutExpired:
            print(name, "TIMEOUT (90 s)")
            continue
        got = [x for x in open(os.path.join(out, name + ".vcf")) if not x.startswith("#")]
        bad = [(a, b) for a, b in zip(ref, got) if a != b]
        stat = [l for l in p.stderr.split("\n") if l.startswith("[sviper_b200] variants=")]
        tr = {}
        for l in p.stderr.split("\n"):
            if l.startswith("[trace] "):
                kv = dict(x.split("=", 1) for x in l.split()[1:])
                tr[kv["refpart"]] = kv
        traces[name] = tr
        print(name, "rc", p.returncode, "records", len(got), "differing", len(bad), "traced", len(tr), stat[-1][:150] if stat else "")
        for a, b in bad[:3]:
            print("   ref:", a.strip()[:200])
            print("   gpu:", b.strip()[:200])
        os.remove(os.path.join(out, name + ".vcf"))
# the device job is a function of its inputs: same (reference part, consensus, reads, statistics) -> same outputs, in every run
INPUTS = ("refpart", "cons", "cons_len", "reads", "reads_hash", "baseQ", "ins")
seen_io, clashes = {}, 0
for nme, tr in traces.items():
    for kv in tr.values():
        key = tuple(kv[f] for f in INPUTS)
        val = tuple(sorted((f, v) for f, v in kv.items() if f not in INPUTS))
        if seen_io.setdefault(key, val) != val:
            clashes += 1
            print("SAME INPUTS, DIFFERENT OUTPUTS:", nme, key, val, seen_io[key])
print("device jobs traced: %d, distinct inputs: %d, same inputs with different outputs: %d" % (sum(len(t) for t in traces.values()), len(seen_io), clashes))
# which fields of which variants differ between runs
names = sorted(traces)
if names:
    first = traces[names[0]]
    for key in first:
        seen = {}
        for nme in names:
            kv = traces[nme].get(key)
            if kv is None:
                continue
            for f, v in kv.items():
                seen.setdefault(f, {}).setdefault(v, []).append(nme)
        for f, vals in seen.items():
            if len(vals) > 1:
                print("variant refpart=%s field %s:" % (key, f))
                for v, who in vals.items():
                    print("    %s  <- %s" % (v[:200], ",".join(w.replace("wp_", "") for w in who)))

"""Writes profiles/r2_summary.md from the committed round-2 evidence: profiles/r2_bench_*.json (bench.py lines),
profiles/r2_configs.json (tools/config_bench.py), profiles/r2_new_paths.json and the ncu captures
profiles/r2_ncu_*.csv (one row per kernel, from `ncu --set full --clock-control none` .ncu-rep files read with
`ncu -i ... --page raw --csv`; tools/r2_prof_all.sh is the capture script). No GPU needed."""
import csv
import glob
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")
out = []
w = out.append


def last_json(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


w("# Round 2 — measured evidence (B200, sm_100a)\n")
w("Everything here was measured on a `gpurun` B200 box with the committed tree; the numbers are CUDA-event times")
w("outside any profiler, the ncu columns come from separate `--set full --clock-control none` captures of the same")
w("commands (`tools/r2_prof_all.sh`). Peaks: `MEASURED_PEAKS.json` (HBM 6544 GB/s; bf16 sustained 1355 TF/s, so dense")
w("TF32 = 678 TF/s).\n")

for name in sorted(glob.glob(os.path.join(P, "r2_bench_*.json"))):
    d = last_json(name)
    if d.get("impl") == "reference":
        c = d["cpu_baseline"]
        w(f"## `{os.path.basename(name)}` — bench.py --impl reference (CPU arm, {c['cores']} host cores)\n")
        w(f"* {d['value']:.0f} vectors/s encoded, {d['adc']['value']/1e6:.0f} M codes/s scanned; {c['sample']}\n")
        continue
    w(f"## `{os.path.basename(name)}` — bench.py, N={d['n_gpus']}\n")
    w(f"* workload: {d['config']['workload']}")
    w(f"* value: **{d['value']/1e6:.1f} M vectors/s** encoded, **{d['adc']['value']/1e9:.1f} G codes/s** ADC "
      f"({d['ms_encode_per_step']:.2f} + {d['ms_scan_per_step']:.2f} ms per step)")
    r = d["roofline"]
    w(f"* roofline ({r['kernel']}): {r['achieved']:.0f} of {r['peak']:.0f} {r['unit']} = **{r['frac']:.2f}** "
      f"(SURVEY 8(d) accounting); DRAM actually moved {r.get('dram_gbs', 0):.0f} GB/s; shared-memory look-up floor "
      f"{r.get('lds_floor_ms', 0):.2f} ms -> lds_frac {r.get('lds_frac', 0):.2f}")
    ra = d.get("roofline_assign")
    if ra:
        w(f"* roofline_assign ({ra['kernel']}): {ra['achieved']:.0f} of {ra['peak']:.0f} TFLOP/s = **{ra['frac']:.2f}**")
    e = d["e2e"]
    w(f"* e2e (host buffers through haag_vq): {e['value']/1e6:.2f} M vectors/s, {e['adc']['value']/1e9:.1f} G codes/s; "
      f"host link {e.get('h2d_gbs_per_rank')} GB/s per rank")
    c = d.get("cpu_baseline", {})
    if c:
        w(f"* cpu_baseline ({c['kind']}, {c['cores']} cores): {c['value']:.0f} vectors/s, "
          f"{c['adc']['value']/1e6:.0f} M codes/s")
    w(f"* parity_check: {d.get('parity_check')}; clocks {d['clocks']['sm_mhz']:.0f}/{d['clocks']['sm_max_mhz']:.0f} MHz "
      f"{d['clocks']['reasons']}; gpu_launches {d['gpu_launches']}\n")

cfg = os.path.join(P, "r2_configs.json")
if os.path.exists(cfg):
    d = json.load(open(cfg))
    ivfj = os.path.join(P, "r2_ivf.json")
    if os.path.exists(ivfj):
        d["rows"] = d["rows"] + json.load(open(ivfj))["rows"]
    w("## Per-config kernel table (`tools/config_bench.py`, BASELINE C1-C4 shapes at full size)\n")
    w("| config | kernel | ms | units/s | GB/s (frac HBM) | TFLOP/s (frac TF32) | note |")
    w("|---|---|---:|---:|---:|---:|---|")
    for r in d["rows"]:
        if "ms" not in r:
            w(f"| {r['config']} | {r['kernel']} | | {r.get('value', 0):.4f} | | | |")
            continue
        gb = f"{r['GBps']:.0f} ({r['frac_hbm']:.2f})" if "GBps" in r else ""
        tf = (f"{r['TFLOPs']:.0f} ({r['frac_tf32']:.2f})" if "frac_tf32" in r else f"{r['TFLOPs']:.0f}") if "TFLOPs" in r else ""
        w(f"| {r['config']} | {r['kernel']} | {r['ms']:.3f} | {r['units_per_s']:.3g} {r['unit']} | {gb} | {tf} | "
          f"{r.get('note', '')} |")
    w("")

npj = os.path.join(P, "r2_new_paths.json")
if os.path.exists(npj):
    d = json.load(open(npj))
    w("## New round-2 paths (`tools/r2_new_paths_bench.py`)\n")
    for k, v in d.items():
        w(f"* `{k}`: {v:.4g}" if isinstance(v, (int, float)) else f"* `{k}`: {v}")
    w("")

w("## ncu captures (`profiles/r2_ncu_*.csv`; per launch, cold cache, under the profiler: shares and pipe")
w("utilisation are the evidence, not the absolute times)\n")
w("| kernel | ms | DRAM MB | tensor % | tmem % | issue % | LSU-shared % | L2 % | regs | grid x block | smem KB |")
w("|---|---:|---:|---:|---:|---:|---:|---:|---:|---|---:|")
for name in sorted(glob.glob(os.path.join(P, "r2_ncu_*.csv"))):
    for r in csv.DictReader(open(name)):
        g = lambda key: next((float(v) for kk, v in r.items() if kk.startswith(key) and v not in ("", None)), 0.0)
        tkey = next(kk for kk in r if kk.startswith('gpu__time_duration'))
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(tkey[tkey.index('[') + 1:tkey.index(']')], 1.0)
        w(f"| {r['kernel']} ({os.path.basename(name)[7:-4]}) | {g('gpu__time_duration') * scale:.3f} | "
          f"{float(r.get('dram_bytes_total') or 0)/1e6:.0f} | {g('sm__pipe_tensor_cycles_active'):.0f} | "
          f"{g('sm__inst_executed_pipe_tmem'):.1f} | {g('smsp__issue_active'):.0f} | "
          f"{g('l1tex__data_pipe_lsu_wavefronts_mem_shared'):.0f} | {g('lts__throughput'):.0f} | "
          f"{g('launch__registers_per_thread'):.0f} | {g('launch__grid_size'):.0f} x {g('launch__block_size'):.0f} | "
          f"{g('launch__shared_mem_per_block_dynamic'):.0f} |")
w("")
w("SASS census of the shipped objects (tcgen05 / TMA / TMEM instruction counts): `profiles/r2_sass_census.txt`.\n")
notes = os.path.join(P, "r2_notes.md")
if os.path.exists(notes):
    w(open(notes).read())
open(os.path.join(P, "r2_summary.md"), "w").write("\n".join(out) + "\n")
print("wrote profiles/r2_summary.md", len(out), "lines")

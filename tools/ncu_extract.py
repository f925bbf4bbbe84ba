"""Turns one `ncu --set full --import-source on` capture of k_step into the text summaries committed under profiles/.

    python tools/ncu_extract.py gpurun_out/r2_final.ncu-rep profiles/r2_ncu_k_step_final "<one-line description>" [n_envs] [ctrl]

Writes  <prefix>_details.txt       (ncu --page details)
        <prefix>_raw_selected.txt  (the raw metrics DESIGN.md cites, tab separated)
        <prefix>_functions.txt     (tools/ncu_funcs.py over the source page)
        <prefix>_lines.txt         (tools/ncu_lines.py, top 60 source lines)
and     profiles/r2_ncu_k_step_traffic.json (dram bytes of the launch + the sha of the kernel sources, read by bench.py)."""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SELECT = re.compile(
    r"^(dram__bytes_(read|write)\.sum(\.per_second|\.pct_of_peak_sustained_elapsed)?|gcc__cache_requests_type_instruction\.sum.*|"
    r"gpu__time_duration\.sum|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum.*|"
    r"launch__(registers_per_thread|shared_mem_per_block_dynamic|grid_size|block_size|occupancy_limit_\w+|waves_per_multiprocessor)|"
    r"sm__icc_request_hit_rate\.pct|sm__icc_requests\.sum.*|sm__inst_executed\.avg\.per_cycle_(elapsed|active)|"
    r"sm__inst_executed_pipe_lsu\.avg\.pct_of_peak_sustained_active|sm__pipe_fma_cycles_active\.avg\.pct_of_peak_sustained_active|"
    r"sm__pipe_alu_cycles_active\.avg\.pct_of_peak_sustained_active|sm__warps_active\.avg\.pct_of_peak_sustained_active|"
    r"sm__issue_active\.avg\.pct_of_peak_sustained_active|sm__inst_issued\.avg\.pct_of_peak_sustained_active|"
    r"smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.sum|smsp__thread_inst_executed\.sum|"
    r"smsp__thread_inst_executed_per_inst_executed\.ratio|smsp__thread_inst_executed_pred_on_per_inst_executed\.ratio|"
    r"smsp__pcsamp_warps_issue_stalled_\w+|smsp__pcsamp_sample_buffer_full|smsp__pcsamp_interval\w*|"
    r"smsp__sass_thread_inst_executed_op_(fadd|fmul|ffma)_pred_on\.sum|sm__sass_inst_executed_op_(shared|local|global)_(ld|st)\.sum|"
    r"smsp__inst_executed_op_(shared|local)_\w+\.sum|lts__t_sector_hit_rate\.pct|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"smsp__sass_thread_inst_executed_op_f(add|mul|fma)_pred_on\.sum\.per_cycle_elapsed|smsp__cycles_elapsed\.avg(\.per_second)?|"
    r"smsp__sass_inst_executed_op_local_(ld|st)\.sum)$")


def ncu(*args):
    return subprocess.run(["ncu", *args], check=True, capture_output=True, text=True).stdout


def main():
    rep, prefix, desc = sys.argv[1], sys.argv[2], sys.argv[3]
    n_envs = int(sys.argv[4]) if len(sys.argv) > 4 else 16384
    ctrl = sys.argv[5] if len(sys.argv) > 5 else "torque"
    open(prefix + "_details.txt", "w").write(ncu("-i", rep, "--page", "details"))
    raw = ncu("-i", rep, "--page", "raw", "--csv")
    rows = list(csv.reader(io.StringIO(raw)))
    names, units, vals = rows[0], rows[1], rows[2]
    sel = {}
    with open(prefix + "_raw_selected.txt", "w") as f:
        f.write(f"# selected raw metrics, ncu --set full --clock-control none --import-source on: {desc}\n")
        for n, u, v in sorted(zip(names, units, vals)):
            if SELECT.match(n):
                f.write(f"{n}\t{u}\t{v}\n"); sel[n] = (u, v)
        tot = sum(float(v.replace(",", "")) for n, (u, v) in sel.items()
                  if n.startswith("smsp__pcsamp_warps_issue_stalled_") and not n.endswith("_not_issued"))
        bar = float(sel["smsp__pcsamp_warps_issue_stalled_barrier"][1].replace(",", ""))
        f.write(f"# derived: barrier share of the stall samples = {bar:.0f} / {tot:.0f} = {bar / tot:.3f}\n")
        g = lambda k: float(sel[k][1].replace(",", ""))
        flop = (g("smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed") + g("smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed")
                + 2 * g("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed")) * g("smsp__cycles_elapsed.avg")
        f.write(f"# derived: executed FP32 work (fadd + fmul + 2 ffma, predicated-on threads) = {flop:.4g} FLOP per launch = {flop / n_envs:.4g} per env-step\n")

    def to_bytes(name):
        u, v = sel[name]; v = float(v.replace(",", ""))
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    from bench import kernel_source_sha
    t = dict(kernel_source_sha=kernel_source_sha(), n_envs=n_envs, ctrl=ctrl, capture=os.path.basename(rep), description=desc,
             dram_bytes_read=to_bytes("dram__bytes_read.sum"), dram_bytes_write=to_bytes("dram__bytes_write.sum"),
             gpu_time_ms_under_ncu=float(sel["gpu__time_duration.sum"][1].replace(",", "")) * {"ms": 1, "us": 1e-3, "s": 1e3}.get(sel["gpu__time_duration.sum"][0], 1))
    t["dram_bytes_per_launch"] = t["dram_bytes_read"] + t["dram_bytes_write"]
    json.dump(t, open(os.path.join(ROOT, "profiles", "r2_ncu_k_step_traffic.json"), "w"), indent=1)
    src = prefix + "_source.csv.tmp"
    open(src, "w").write(ncu("-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"))
    c = os.path.join(ROOT, "quadruped_rl_locomotion_b200", "csrc")
    files = [os.path.join(c, f) for f in ("go1_warp.cuh", "go1_pair.cuh", "go1sim.cu")]
    open(prefix + "_functions.txt", "w").write(
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_funcs.py"), src, *files], check=True, capture_output=True, text=True).stdout)
    open(prefix + "_lines.txt", "w").write(
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), src, "60"], check=True, capture_output=True, text=True).stdout)
    os.remove(src)
    print(json.dumps(t))


if __name__ == "__main__":
    main()

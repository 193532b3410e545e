#!/bin/bash
# FP32 operations per shaded vertex (SURVEY 8d "algorithmic FLOPs per bounce"): executed FADD / FMUL / FFMA thread instructions
# of every k_shade launch of a fixed workload (ncu), divided by the vertices those launches shaded (device counter).
mkdir -p gpurun_out
M=smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__inst_executed_pipe_xu.sum,smsp__thread_inst_executed.sum
for sc in cornell gallery; do
  python scripts/shade_ops_probe.py $sc > gpurun_out/shade_ops_${sc}_plain.json || exit 1
  timeout 900 ncu --clock-control none -k regex:"k_shade" --csv --metrics $M --log-file gpurun_out/shade_ops_${sc}.csv python scripts/shade_ops_probe.py $sc > gpurun_out/shade_ops_${sc}_ncu.json 2> gpurun_out/shade_ops_${sc}.err
done
python - <<'PY'
import csv, json
out = {"source": "ncu smsp__sass_thread_inst_executed_op_{fadd,fmul,ffma}_pred_on over every k_shade launch of 640x360 x 16 spp, depth 8 "
                 "(scripts/gpu_r2_shade_ops.sh), divided by gsb_counters.shaded_vertices of the same run; FFMA counts 2", "scenes": {}}
for sc in ("cornell", "gallery"):
    v = json.loads(open(f"gpurun_out/shade_ops_{sc}_ncu.json").read().strip().splitlines()[-1])
    tot, per_type = {}, {}
    for r in csv.reader(open(f"gpurun_out/shade_ops_{sc}.csv")):
        if len(r) > 14 and r[0].isdigit():
            tot[r[12]] = tot.get(r[12], 0.0) + float(r[14].replace(",", ""))
            k = r[4].split("k_shade<")[1].split(",")[0]
            per_type.setdefault(k, {})
            per_type[k][r[12]] = per_type[k].get(r[12], 0.0) + float(r[14].replace(",", ""))
    f = lambda t: t.get("smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", 0) + t.get("smsp__sass_thread_inst_executed_op_fmul_pred_on.sum", 0) + 2 * t.get("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", 0)
    out["scenes"][sc] = {"shaded_vertices": v["shaded_vertices"], "flops": f(tot), "flops_per_vertex": f(tot) / v["shaded_vertices"],
                         "thread_instructions_per_vertex": tot.get("smsp__thread_inst_executed.sum", 0) / v["shaded_vertices"],
                         "flops_by_bsdf_type": {k: f(t) for k, t in per_type.items()}}
out["flops_per_vertex"] = out["scenes"]["cornell"]["flops_per_vertex"]
json.dump(out, open("gpurun_out/shade_ops.json", "w"), indent=1)
print(json.dumps(out, indent=1))
PY

"""profiles/<name>.json for bench.py's roofline objects from an ncu --set full capture of env_step_kernel at the pinned state:
python tools/ncu_to_json.py <rep> <num_envs> <out.json>.  DRAM traffic and executed FP32 thread operations of the launch come from
the capture, the algorithmic FLOPs per env step from profiles/r02_oracle_flops.json (tools/count_oracle_flops.py)."""
import csv
import io
import json
import os
import subprocess
import sys

rep, num_envs, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, val = rows[0], rows[1], rows[2]


def get(name):
    i = hdr.index(name)
    v, u = float(val[i].replace(",", "")), units[i].lower()
    scale = {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "byte": 1.0, "ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "msecond": 1.0, "ms": 1.0, "nsecond": 1e-6, "inst": 1.0}
    return v * scale.get(u, 1.0)


ffma, fadd, fmul = (get(f"smsp__sass_thread_inst_executed_op_{k}_pred_on.sum") for k in ("ffma", "fadd", "fmul"))
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
orc = json.load(open(os.path.join(root, "profiles", "r02_oracle_flops.json")))["tasks"]["Humanoid-v4"]
d = {"kernel": "env_step_kernel<CfgHumanoid, 8, lockstep>", "num_envs": num_envs, "state": "tests/golden/walking_Humanoid-v4.npz, 8 warm rollouts",
     "kernel_ms_ncu": get("gpu__time_duration.sum"), "dram_bytes_per_launch": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"),
     "executed_fp32_flop_per_launch": 2 * ffma + fadd + fmul, "executed_warp_instructions": get("smsp__inst_executed.sum"),
     "oracle_flop_per_env_step": orc["flop_per_env_step"], "oracle_flop_per_forward_evaluation": orc["flop_per_forward_evaluation"],
     "source": f"ncu --set full capture {os.path.basename(rep)} (executed: 2 x FFMA + FADD + FMUL thread instructions); algorithmic FLOPs counted by the FP64 "
               "oracle with an op-counting scalar over the 256 walking states (tools/count_oracle_flops.py)"}
json.dump(d, open(out, "w"), indent=1)
print(json.dumps(d))

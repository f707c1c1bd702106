"""DRAM traffic per launch (dram__bytes_read.sum + dram__bytes_write.sum) of each kernel in an
`ncu --set full` report, keyed by bench.py's kernel names -> profiles/<round>_ncu_traffic.json."""
import csv, json, subprocess, sys, collections
NAMES = [
    ("tc_lstm_fwd_kernel", "tc_lstm_fwd_kernel"), ("tc_lstm_bwd_kernel", "tc_lstm_bwd_kernel"), ("tc_lstm_bwd_pair_kernel", "tc_lstm_bwd_kernel"),
    ("tc_conv1_sw_dual_kernel", "tc_conv1_fwd_dual"), ("tc_conv1_wgrad_sw_kernel", "tc_conv1_wgrad"),
    ("tc_conv2_wgrad_sw_kernel", "tc_conv2_wgrad"), ("tc_conv3_wgrad_sw_kernel", "tc_conv3_wgrad"),
    ("tc_conv2_dgrad_sw_kernel", "tc_conv2_dgrad"), ("tc_conv3_dgrad_sw_kernel", "tc_conv3_dgrad"),
    ("tc_gemm_x3_pair_kernel", "tc_gemm_kernel_x3"), ("tc_gemm_x3_kernel", "tc_gemm_kernel_x3"),
    ("tc_gemm_tn_kernel", "tc_gemm_tn_kernel"), ("tc_gemm_nn_kernel", "tc_gemm_nn_kernel"),
    ("rmsprop_kernel", "rmsprop_kernel"),
    ("tc_conv_fwd_kernel<unsigned char", "tc_conv1_fwd"), ("tc_conv_fwd_kernel<float, 32, 4", "tc_conv2_fwd"),
    ("tc_conv_fwd_kernel<float, 32, 3", "tc_conv3_fwd"),
    ("tc_conv_sw_fwd_kernel<2>", "tc_conv2_fwd"), ("tc_conv_sw_fwd_kernel<3>", "tc_conv3_fwd"),
    ("tc_conv_wgrad_kernel<unsigned char", "tc_conv1_wgrad"), ("tc_conv_wgrad_kernel<float, 32, 4", "tc_conv2_wgrad"),
    ("tc_conv_wgrad_kernel<float, 32, 3", "tc_conv3_wgrad"),
    ("tc_conv_dgrad_kernel<4", "tc_conv2_dgrad"), ("tc_conv_dgrad_kernel<3", "tc_conv3_dgrad"),
    ("tc_conv_dgrad2_kernel", "tc_conv2_dgrad"),
    ("tc_gemm_kernel", "tc_gemm_kernel*"),
]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
ik, ir, iw, it = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for r in rows[2:]:
    for pat, key in NAMES:
        if pat in r[ik]:
            a = agg[key]
            a[0] += 1
            a[1] += float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
            a[2] += float(r[it])
            break
res = {k: {"launches": n, "dram_bytes_per_launch": b / n, "avg_duration_" + units[it]: t / n} for k, (n, b, t) in agg.items()}
# provenance: bench.py reports `roofline.traffic` from this file only while csrc/ is unchanged
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import csrc_sha16
try:
    commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
except OSError:
    commit = ""
res["_meta"] = {"csrc_sha16": csrc_sha16(), "commit": commit or (sys.argv[3] if len(sys.argv) > 3 else None),
                "source": sys.argv[1] + " (ncu --set full --clock-control none, one learner step, "
                          "B=64 T=40+40 n=5 H=512)"}
json.dump(res, open(sys.argv[2], "w"), indent=1)
print(json.dumps(res, indent=1))

"""Copy the summaries of one tools/gpu_r2_final.sh pass from gpurun_out/ into profiles/ (tracked) and rebuild
profiles/ncu_traffic.json in the format bench.py reads.   python tools/collect_profiles.py r02_f1 r02"""
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src_tag, dst_tag = sys.argv[1], sys.argv[2]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def cp(a, b):
    if os.path.exists(os.path.join(G, a)):
        shutil.copy(os.path.join(G, a), os.path.join(P, b))
        return True
    print("missing", a)
    return False


for wl in ("c1_poisson_finetune_3x64_1k", "c2_poisson_leap_3x64_T16_1k", "c3_burgers_maml_3x64_T8_1k", "c4_hyper_maml_3x64_T8_1k"):
    cp(f"{src_tag}_bench_{wl}.json", f"{dst_tag}_bench_{wl}.json")
cp(f"{src_tag}_bench_c5.json", f"{dst_tag}_bench_c5.json")
cp(f"{src_tag}_bench_c5_reference_arm.json", f"{dst_tag}_bench_c5_reference_arm.json")
cp(f"{src_tag}_pytest.log", f"{dst_tag}_gpu_pytest.log")
cp(f"{src_tag}_smoke.log", f"{dst_tag}_smoke.log")
cp(f"{src_tag}_ncu_launches_c5.md", f"{dst_tag}_ncu_launches_c5.md")
cp(f"{src_tag}_deploy_time.jsonl", f"{dst_tag}_deploy_time.jsonl")
for mode, kn in (("hvp", "tc_hvp_kernel"), ("grad", "tc_pde_kernel"), ("loss", "tc_pde_kernel"), ("grad", "tc_wgrad_kernel"), ("hvp", "tc_wgrad_kernel")):
    cp(f"{src_tag}_ncu_{mode}_{kn}_details.txt", f"{dst_tag}_ncu_full_{kn}_{mode}_details.txt")
    cp(f"{src_tag}_ncu_{mode}_{kn}_source_summary.txt", f"{dst_tag}_ncu_full_{kn}_{mode}_sass_stalls.txt")
tj = json.load(open(os.path.join(G, f"{src_tag}_ncu_traffic.json")))
out = {"workload": "c5_poisson_maml_5x256_T64_16k",
       "how": f"ncu --set full --clock-control none --import-source on, one launch each at C5 size (tools/gpu_r2_final.sh {src_tag}: "
              "tools/ncu_one.py hvp|grad|loss on one B200); traffic_bytes = dram__bytes_read.sum + dram__bytes_write.sum per launch",
       "kernels": {"tc_hvp_kernel": tj.get("hvp_tc_hvp_kernel"), "tc_pde_kernel_grad": tj.get("grad_tc_pde_kernel"),
                   "tc_pde_kernel_loss": tj.get("loss_tc_pde_kernel"), "tc_wgrad_kernel_grad": tj.get("grad_tc_wgrad_kernel"),
                   "tc_wgrad_kernel_hv": tj.get("hvp_tc_wgrad_kernel")}}
json.dump(out, open(os.path.join(P, "ncu_traffic.json"), "w"), indent=1)
print("ok")

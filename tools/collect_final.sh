#!/bin/bash
# Copies what tools/gpu_final.sh left under gpurun_out/ into the tracked profiles/ files of round 2 (run from the repo root).
O=gpurun_out
grep '^{' $O/final_bench.json | tail -1 > profiles/r02_bench_b256_final.json
grep '^{' $O/final_reference.json | tail -1 > profiles/r02_bench_reference_arm_final.json
{ echo "# python -m pytest tests -m gpu -q -s   (one B200, final round-2 build; tools/gpu_final.sh)"
  grep -v "Warning\|warnings.warn\|^$\|register_buffer\|^  /root\|^    \|DDIM Sampler\|^tests/test_" $O/final_pytest.log | cut -c1-420
  echo; echo "# python -c 'import __graft_entry__ as g; g.smoke()'"; tail -2 $O/final_smoke.log; } > profiles/r02_gpu_tests_final.txt
{ echo "# ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none  python tools/one_step.py train"
  echo "# ONE eager UNet training step at B = 256 (the step bench.py captures), final round-2 build; times are cold-cache and serialised"
  cat $O/final_dram_summary.txt; echo; cat $O/final_launches.txt; } > profiles/r02_step_launches_final.txt
{ echo "# ncu ... python tools/one_step.py full : ONE eager full step from images (EncDiffTrainer.step: VQ encode -> concept encoder -> UNet step -> encoder backward + AdamW), B = 256"
  python tools/ncu_dram_summary.py $O/final_launches_full.csv /tmp/x.json; echo; cat $O/final_launches_full.txt; } > profiles/r02_full_step_launches_final.txt
cp $O/final_sgemm2_dram.json profiles/r02_sgemm2_dram.json
cp $O/final_ablation.txt profiles/experiments/r02_ablation_final.txt
grep '^{' $O/final_vq_bench.json | tail -1 > profiles/r02_vq_first_stage_bench.json
{ echo "# Concept-encoder trunk (encdiff_enc4_*) and the full step from images, one B200, final round-2 build"
  echo "## tests (from profiles/r02_gpu_tests_final.txt)"
  grep "enc4 trunk\|Encoder4 drop-in\|full step from images\|targets swapped\|full pipeline" profiles/r02_gpu_tests_final.txt | cut -c1-400
  echo; echo "## python tools/enc4_profile.py  (B = 256; first parity-green build: 1.617 ms, before the loads-in-flight pass over the elementwise kernels)"
  cat $O/final_enc4_profile.txt; echo
  echo "## python tools/exp_pipeline_breakdown.py  (B = 256, CUDA-event time per phase of the full step; 'CUDA trunk' is the product path,"
  echo "## the 'torch trunk' lines run Encoder4(torch_trunk=True), i.e. the reference's layer sequence in PyTorch on the same GPU)"
  cat $O/final_pipeline_breakdown.txt; echo
  echo "## bench.py full_pipeline (EncDiffTrainer.step from pinned host images, B = 256)"
  python -c "import json; d=json.load(open('profiles/r02_bench_b256_final.json')); print(json.dumps(d['full_pipeline']))"; } > profiles/r02_enc4_trunk.txt
python - <<'PY'
import json
d=json.load(open('profiles/r02_bench_b256_final.json'))
print("step", d['ms_per_step'], d['value'], "e2e", d['e2e']['value'], "frac", d['roofline']['frac'], d['roofline']['achieved'], "eager", d['roofline']['eager_events']['frac'], d['roofline']['eager_events']['achieved'], "step_frac", d['step_tensor_frac'])
print("ddim", d['ddim50']['img_per_s'], "swap", d['swap_ddim200']['img_per_s'], "full", d['full_pipeline']['ms_per_step'], d['full_pipeline']['img_per_s'], "cpu", d['cpu_baseline']['value'], d['config1_cpu_full_step']['img_per_s'])
c=d['comparators']; print(c['ref_eager_gpu_tf32'], c['ref_eager_gpu_bf16']['train_img_per_s'], c['dropin_autograd']['train_ms_per_step'], c['dropin_autograd']['train_img_per_s'])
print("ref arm", json.load(open('profiles/r02_bench_reference_arm_final.json'))['value'])
v=json.load(open('profiles/r02_vq_first_stage_bench.json')); print("vq", v['value'], v['ms_per_batch_graph'], v['decode']['value'], v['decode']['ms_per_batch_graph'])
PY
grep passed profiles/r02_gpu_tests_final.txt; head -4 gpurun_out/final_pipeline_breakdown.txt | cut -c1-200; grep "sgemm2 all\|wgrad2_group\|gn_fwd\|gn_bwd\|ln_fwd\|ln_bwd\|geglu\|self_attn_fwd \|self_attn_bwd \|all norms\|all attention\|^(none)" gpurun_out/final_ablation.txt

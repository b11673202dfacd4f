# ncu captures behind profiles/r2_ncu_*.csv (run on the GPU box: bash tools/r2_prof_all.sh). Each workload first runs
# plain (must exit 0), then its kernels are captured one by one under `ncu --set full --clock-control none`; each
# .ncu-rep is summarised on the spot with tools/ncu_summary.py (gpurun_out/r2_ncu_<name>.csv) and removed (gpurun
# carries at most 64 MiB back).
set -x
declare -A OK
P() { # name regex skip count cmd...
  name=$1; rx=$2; sk=$3; ct=$4; shift 4
  key="$*"
  if [ -z "${OK[$key]}" ]; then "$@" > gpurun_out/plain_$name.log 2>&1 || { echo "$name plain run failed"; return; }; OK[$key]=1; fi
  ncu --set full --clock-control none -k regex:$rx -s $sk -c $ct -o gpurun_out/r2_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  rc=$?
  python tools/ncu_summary.py gpurun_out/r2_$name.ncu-rep gpurun_out/r2_ncu_$name.csv && rm -f gpurun_out/r2_$name.ncu-rep
  echo "$name rc=$rc"
}
P scan_screen adc_screen_kernel 1 1 python tools/r2_prof.py scan
P scan_finish screen_finish_kernel 1 1 python tools/r2_prof.py scan
P rotate rotate_tc_kernel 1 1 python tools/r2_prof.py rotate
P caq caq_encode_t_segments_kernel 1 1 python tools/r2_prof.py caq
P flat_screen flat_screen_kernel 4 4 python tools/flat_prof.py
P flat_pool flat_pool_kernel 3 3 python tools/flat_prof.py
P flat_finish flat_finish_kernel 1 1 python tools/flat_prof.py
P rq_screen rq_screen_kernel 11 1 python tools/rq_prof.py
P assign assign_tc_kernel 2 1 python tools/assign_prof.py
P est_screen est_screen_kernel 4 4 python tools/r2_prof_est.py
P est_pool est_pool_kernel 3 3 python tools/r2_prof_est.py
P erq_pack erabitq_pack_tc 1 1 python tools/r2_prof_est.py
P erq_fixup erabitq_fixup_kernel 1 1 python tools/r2_prof_est.py
P lvq_encode lvq_encode2p_kernel 3 1 python tools/config_bench.py c3 gpurun_out/_c3.json
P lvq_decode lvq_decode4_kernel 3 1 python tools/config_bench.py c3 gpurun_out/_c3.json
P saq_encode saq_encode_vec_kernel 3 1 python tools/config_bench.py c3 gpurun_out/_c3.json
P rq1_encode rabitq1_encode_vec_kernel 3 1 python tools/config_bench.py c4 gpurun_out/_c4.json
P c1_lut adc_lut_batched_kernel 3 1 python tools/r2_prof.py c1adc
P c1_sample adc_sample_kth_kernel 3 1 python tools/r2_prof.py c1adc
P c1_screen adc_screen_kernel 3 1 python tools/r2_prof.py c1adc
P c1_finish screen_finish_kernel 3 1 python tools/r2_prof.py c1adc
rm -f gpurun_out/*.ncu-rep
du -sh gpurun_out

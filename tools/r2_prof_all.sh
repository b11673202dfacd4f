set -x
P() { # name regex skip count cmd...
  name=$1; rx=$2; sk=$3; ct=$4; shift 4
  "$@" > gpurun_out/plain_$name.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$rx -s $sk -c $ct -o gpurun_out/r2_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
}
P scan_screen adc_screen_kernel 1 1 python tools/r2_prof.py scan
P scan_finish screen_finish_kernel 1 1 python tools/r2_prof.py scan
P rotate rotate_tc_kernel 1 1 python tools/r2_prof.py rotate
P caq caq_encode_t_segments_kernel 1 1 python tools/r2_prof.py caq
P flat_screen flat_screen_kernel 12 2 python tools/flat_prof.py
P flat_finish flat_finish_kernel 14 1 python tools/flat_prof.py
P rq_screen rq_screen_kernel 11 1 python tools/rq_prof.py
P assign assign_tc_kernel 2 1 python tools/assign_prof.py
P est_screen est_screen_kernel 5 1 python tools/r2_prof_est.py
P erq_pack erabitq_pack_tc 1 1 python tools/r2_prof_est.py
ls -la gpurun_out/*.ncu-rep

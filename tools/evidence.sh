#!/bin/bash
# One-GPU evidence run for profiles/: tests, smoke, the bench lines (never under a profiler), then the ncu passes of
# B200_PROFILING.md (launch list + one --set full capture per top kernel). Usage on the GPU box:
#   bash tools/evidence.sh TAG        -> gpurun_out/TAG_*
# Each ncu pass runs only after the same command exited 0 without ncu.
set -u
T=${1:-ev}
O=gpurun_out
mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on"

timeout 900 python -m pytest tests -m gpu -q > $O/${T}_tests.log 2>&1; echo "tests rc=$?" | tee -a $O/${T}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${T}_smoke.log
timeout 900 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_ref.json 2> $O/${T}_ref.err; echo "ref rc=$?"

if timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > /dev/null 2> $O/${T}_pre.err; then
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${T}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu > $O/${T}_ncu.log 2>&1; echo "launch list rc=$?"
fi
if timeout 300 python tools/prof_target.py large 64 8 > /dev/null 2>&1; then
  timeout 900 $NCU -k regex:large_update -c 3 -o $O/${T}_large -f python tools/prof_target.py large 64 8 > $O/${T}_ncu_large.log 2>&1; echo "large rc=$?"
fi
if timeout 300 python tools/prof_target.py batched 65536 8 > /dev/null 2>&1; then
  timeout 900 $NCU -k regex:ekf_batched -s 1 -c 2 -o $O/${T}_batched -f python tools/prof_target.py batched 65536 8 > $O/${T}_ncu_b.log 2>&1; echo "batched rc=$?"
fi
if timeout 300 python tools/prof_target.py laser 65536 > /dev/null 2>&1; then
  timeout 900 $NCU -k 'regex:laser_(extract|fit)' -s 2 -c 4 -o $O/${T}_laser -f python tools/prof_target.py laser 65536 > $O/${T}_ncu_laser.log 2>&1; echo "laser rc=$?"
fi
# round 2: the solve kernel of the cfg4 step, the cfg5 kernels (DMMA GEMM, Cholesky panel, 256-block inverse; TF32 GEMM), the batched UKF
if timeout 300 python tools/prof_target.py large 64 8 > /dev/null 2>&1; then
  timeout 900 $NCU -k regex:large_solve -s 3 -c 1 -o $O/${T}_solve -f python tools/prof_target.py large 64 8 > $O/${T}_ncu_solve.log 2>&1; echo "solve rc=$?"
fi
if timeout 300 python tools/prof_target.py ukf_large 32 > /dev/null 2>&1; then
  timeout 900 $NCU -k regex:dgemm_nt_kernel -s 150 -c 6 -o $O/${T}_dgemm -f python tools/prof_target.py ukf_large 32 > $O/${T}_ncu_dgemm.log 2>&1; echo "dgemm rc=$?"
  timeout 900 $NCU -k 'regex:chol_panel|tri_inv256' -s 30 -c 3 -o $O/${T}_chol -f python tools/prof_target.py ukf_large 32 > $O/${T}_ncu_chol.log 2>&1; echo "chol rc=$?"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/${T}_launches_ukf_large.csv python tools/prof_target.py ukf_large 32 > $O/${T}_ncu_ukfl.log 2>&1; echo "ukf_large launch list rc=$?"
fi
if timeout 300 python tools/prof_target.py ukf_large 32 1 > /dev/null 2>&1; then
  timeout 900 $NCU -k regex:tf32_gemm -s 3 -c 3 -o $O/${T}_tf32 -f python tools/prof_target.py ukf_large 32 1 > $O/${T}_ncu_tf32.log 2>&1; echo "tf32 rc=$?"
fi
if timeout 300 python tools/prof_target.py ukf 65536 8 > /dev/null 2>&1; then
  timeout 900 $NCU -k regex:ukf_batched -s 1 -c 1 -o $O/${T}_ukfb -f python tools/prof_target.py ukf 65536 8 > $O/${T}_ncu_ukfb.log 2>&1; echo "ukf batched rc=$?"
fi
# gpurun brings back at most 64 MiB: keep the raw-page CSV of every capture (what tools/collect_profiles.py reads) and
# drop the .ncu-rep files unless KEEP_REP=1
for rep in $O/${T}_*.ncu-rep; do
  [ -f "$rep" ] || continue
  ncu -i "$rep" --page raw --csv > "${rep%.ncu-rep}.raw.csv" 2>/dev/null
  [ "${KEEP_REP:-0}" = "1" ] || rm -f "$rep"
done
ls -la $O | grep ${T}_

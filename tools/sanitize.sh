#!/bin/bash
# compute-sanitizer passes over the hand-rolled synchronisation of the library (SURVEY section 5 "race detection / sanitizers"):
#   memcheck  : out-of-bounds / misaligned accesses (the compacted terminal rows, shared staging tiles, cp.async slots)
#   racecheck : shared-memory hazards (k_vstep's warp-private lists and staging tile, vstep_flush, k_step's block-level compaction,
#               k_emm_fused's tables, k_playouts' pick-up path)
#   synccheck : barrier misuse (__syncthreads / __syncwarp in divergent code)
# over (a) __graft_entry__.smoke(): K1 persistent + block kernels, K3, K4 fused; (b) the single-process peer-merge tests (flag words,
# epochs, time-out path of k_mc_best_peers) and the VecEnv contract tests; (c) with >= 2 GPUs the two-process NVLink merge.
# Run on the GPU box:  bash tools/sanitize.sh [outdir]      Logs: <outdir>/sanitize_<tool>_<what>.log, summary on stdout.
set -u
cd "$(dirname "$0")/.."
out=${1:-gpurun_out}
mkdir -p "$out"
CS=${COMPUTE_SANITIZER:-compute-sanitizer}
ngpu=$(python -c "import torch; print(torch.cuda.device_count())" 2>/dev/null || echo 0)
status=0
run() {   # run <tool> <what> <command...>
    tool=$1; what=$2; shift 2
    log="$out/sanitize_${tool}_${what}.log"
    timeout 1500 $CS --tool "$tool" --print-limit 20 --error-exitcode 86 "$@" > "$log" 2>&1
    rc=$?
    tail_line=$(grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" "$log" | tail -n 2 | tr '\n' ' ')
    echo "[$tool] $what: rc=$rc  $tail_line"
    [ $rc -ne 0 ] && status=1
}
SMOKE=(python -c "import __graft_entry__ as g; g.smoke()")
PEER=(python -m pytest -x -q -m gpu tests/test_peer.py -k "fused_merge or times_out" -p no:cacheprovider)
VEC=(python -m pytest -x -q -m gpu tests/test_vec_contract.py tests/test_vec_env.py -k "outputs_of_a_step or persistent_step_kernel_vs_block or in_kernel_random_agent or messages_come" -p no:cacheprovider)
SEARCH=(python -m pytest -x -q -m gpu tests/test_engine_parity.py tests/test_sim_winrate.py tests/test_uct.py -k "fused_equals_replay or cuda_kernels_equal or counters_equal" -p no:cacheprovider)
for tool in memcheck racecheck synccheck; do
    run $tool smoke "${SMOKE[@]}"
    run $tool peer "${PEER[@]}"
    run $tool vec "${VEC[@]}"
    run $tool search "${SEARCH[@]}"
done
if [ "$ngpu" -ge 2 ]; then
    run memcheck peer2 python -m pytest -x -q -m gpu tests/test_peer.py -k "processes_merge_over_nvlink and 2" -p no:cacheprovider
fi
exit $status

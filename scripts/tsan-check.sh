#!/bin/sh
# ThreadSanitizer pass over the host runtime: builds every C++ source of libkungfu_host together with
# the C++ agent test (tests/cpp/fake_agent.cpp: all collectives, async ops with callbacks, p2p store,
# consensus, AllGatherTransform) with -fsanitize=thread and runs it as a 2- and a 3-worker job.
# The reference has no sanitizer run at all. Exit code != 0 on any TSAN report.
set -e
cd "$(dirname "$0")/.."
out=${TMPDIR:-/tmp}/kungfu-tsan
mkdir -p "$out"
srcs=$(ls csrc/core/*.cpp csrc/ctrl/*.cpp csrc/elastic/*.cpp csrc/capi/*.cpp)
g++ -O1 -g -std=c++17 -fsanitize=thread -fPIE -pie -pthread -mavx2 -mf16c -Icsrc -Iinclude \
    $srcs tests/cpp/fake_agent.cpp -o "$out/fake-agent-tsan" -ldl
status=0
for cfg in "2 STAR" "3 BINARY_TREE_STAR" "3 RING"; do
    set -- $cfg
    base=$((24000 + $$ % 4000))
    log="$out/tsan-np$1-$2.log"
    TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" timeout 300 kungfu_b200/bin/kungfu-run -q -np "$1" -strategy "$2" \
        -timeout 250s -port-range "$base-$((base + 16))" -port "$((base + 20))" "$out/fake-agent-tsan" >"$log" 2>&1 || status=1
    reports=$(grep -c "WARNING: ThreadSanitizer" "$log" || true)
    ok=$(grep -c "OK fake-agent" "$log" || true)
    echo "np=$1 strategy=$2: $ok/$1 workers ok, $reports TSAN reports ($log)"
    [ "$reports" = "0" ] && [ "$ok" = "$1" ] || status=1
done
# elastic paths: an instrumented trainer that resizes 1 -> 3 -> 4 -> 2 under a normal watch-mode runner, then
# a normal trainer under an instrumented runner
g++ -O1 -g -std=c++17 -fsanitize=thread -fPIE -pie -pthread -mavx2 -mf16c -Icsrc -Iinclude \
    $srcs tests/cpp/fake_adaptive_trainer.cpp -o "$out/adaptive-tsan" -ldl
g++ -O1 -g -std=c++17 -fsanitize=thread -fPIE -pie -pthread -mavx2 -mf16c -Icsrc -Iinclude \
    $srcs csrc/launch/kungfu_run_main.cpp -o "$out/kungfu-run-tsan" -ldl
g++ -O1 -std=c++17 -pthread -Iinclude tests/cpp/fake_adaptive_trainer.cpp -o "$out/adaptive" \
    -Lkungfu_b200/lib -lkungfu_host -Wl,-rpath,"$PWD/kungfu_b200/lib"
for pair in "kungfu_b200/bin/kungfu-run $out/adaptive-tsan worker" "$out/kungfu-run-tsan $out/adaptive runner"; do
    set -- $pair
    base=$((26000 + $$ % 1500))
    log="$out/tsan-elastic-$3.log"
    TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" timeout 300 "$1" -q -w -np 1 -H 127.0.0.1:4 -timeout 250s \
        -port-range "$base-$((base + 32))" -port "$((base + 40))" "$2" 3:1,3:3,3:4,3:2 20000 >"$log" 2>&1 || status=1
    reports=$(grep -c "WARNING: ThreadSanitizer" "$log" || true)
    ok=$(grep -c "OK fake-adaptive-trainer" "$log" || true)
    echo "elastic 1->3->4->2, instrumented $3: $ok/2 survivors ok, $reports TSAN reports ($log)"
    [ "$reports" = "0" ] && [ "$ok" = "2" ] || status=1
done
# AddressSanitizer + UndefinedBehaviorSanitizer + leak check over the same program
g++ -O1 -g -std=c++17 -fsanitize=address,undefined -fno-omit-frame-pointer -pthread -mavx2 -mf16c -Icsrc -Iinclude \
    $srcs tests/cpp/fake_agent.cpp -o "$out/fake-agent-asan" -ldl
base=$((28000 + $$ % 2000))
log="$out/asan-np3.log"
ASAN_OPTIONS="detect_leaks=1:halt_on_error=0" UBSAN_OPTIONS="print_stacktrace=1" timeout 300 kungfu_b200/bin/kungfu-run -q -np 3 \
    -strategy MULTI_BINARY_TREE_STAR -timeout 250s -port-range "$base-$((base + 16))" -port "$((base + 20))" \
    "$out/fake-agent-asan" >"$log" 2>&1 || status=1
reports=$(grep -c -E "ERROR: AddressSanitizer|runtime error|LeakSanitizer" "$log" || true)
ok=$(grep -c "OK fake-agent" "$log" || true)
echo "asan+ubsan np=3: $ok/3 workers ok, $reports reports ($log)"
[ "$reports" = "0" ] && [ "$ok" = "3" ] || status=1
exit $status

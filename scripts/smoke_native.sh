#!/bin/bash
# Ten-second check of the native stack on one GPU without Python: daemon with an HBM pool and a VMM host pool,
# ocm_test selectors on GPU and host kinds, both allocation paths, unit allocator suites.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
NF=/tmp/ocm_smoke_nodefile
echo "0 localhost 127.0.0.1 19555 0 gpu=0 gpu_bytes=1G host_bytes=256M" > $NF
export OCM_NS=smk$$ OCM_LOG=info OCM_RANK=0
timeout 60 build/bin/oncillamem $NF > gpurun_out/smoke_daemon.log 2>&1 &
D=$!
for i in $(seq 1 100); do grep -q "oncillamem ready" gpurun_out/smoke_daemon.log && break; sleep 0.1; done
rc=0
for args in "1 1 2 7" "2 1 1 7" "3 1 1 7" "2 1 1 3" "3 1 1 1" "5 1 1 8"; do
  timeout 20 build/bin/ocm_test $args > gpurun_out/smoke_t.log 2>&1 || { rc=1; echo "FAIL ocm_test $args"; tail -3 gpurun_out/smoke_t.log; }
  OCM_ALLOC_PATH=daemon timeout 20 build/bin/ocm_test $args > gpurun_out/smoke_t.log 2>&1 || { rc=1; echo "FAIL (daemon path) ocm_test $args"; tail -3 gpurun_out/smoke_t.log; }
done
timeout 20 build/bin/hello_ocm > gpurun_out/smoke_hello.log 2>&1 || { rc=1; echo "FAIL hello_ocm"; tail -3 gpurun_out/smoke_hello.log; }
kill -TERM $D; wait $D 2>/dev/null
grep -h "pools:\|ocm E\|ocm W" gpurun_out/smoke_daemon.log | head -5
echo "smoke_native rc=$rc"
exit $rc

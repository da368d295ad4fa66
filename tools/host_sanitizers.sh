# Host library under the sanitizers (CPU only; the reference's CI runs its unit tests under ASAN + UBSAN).
#   bash tools/host_sanitizers.sh      # rebuilds the normal library at the end
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
cd "$ROOT/gelex_b200/csrc/host"
SRCS="plink_meta.cpp grm_engine.cpp grm_loader.cpp genotype_pipe.cpp capi.cpp"
# 1. ASAN + UBSAN: the CPU tests of the host side (metadata, plan, writers, relay protocol, differential tests against the reference)
g++ -std=c++20 -O1 -g -fPIC -pthread -fsanitize=address,undefined -fno-omit-frame-pointer -shared -o "$ROOT/gelex_b200/libggrm_host.so" $SRCS \
    -L"$ROOT/gelex_b200" -lggrm_b200 -Wl,-rpath,'$ORIGIN'
cd "$ROOT"
LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)" ASAN_OPTIONS=detect_leaks=0 \
    python -m pytest tests/test_host_engine.py tests/test_reference_differential.py tests/test_sharding.py -x -q -m "not gpu" -p no:cacheprovider -s > /tmp/host_asan.log 2>&1 || true
echo "ASAN/UBSAN reports: $(grep -c 'runtime error\|AddressSanitizer' /tmp/host_asan.log)"; tail -1 /tmp/host_asan.log
# 2. TSAN: the two multi-threaded pieces that run without a GPU (ChunkRelay = BED relay protocol, WriterPool + BinFile)
cat > /tmp/tsan_main.cpp <<'CPP'
#include <cstdio>
#include "grm_engine.h"
using namespace gelex::b200;
int main() {
    long bad = 0;
    for (int w : {1, 3, 8}) bad += GrmEngine::relay_selftest(400, w, 2, 4096);
    try { GrmEngine::relay_selftest(200, 4, 2, 512, 37); } catch (const std::exception& e) { std::printf("expected: %s\n", e.what()); }
    bool mapped = false;
    for (int tiles = 0; tiles < 2; ++tiles)
        for (auto mode : {WriterMode::Pwrite, WriterMode::Mmap})
            for (int prefault : {0, 3}) GrmEngine::writer_selftest("/dev/shm/tsan_w.bin", 4500, tiles != 0, 3, 2, mode, &mapped, prefault);
    std::printf("bad=%ld\n", bad);
    return bad != 0;
}
CPP
cd "$ROOT/gelex_b200/csrc/host"
g++ -std=c++20 -O1 -g -pthread -fsanitize=thread -I. -o /tmp/tsan_host /tmp/tsan_main.cpp $SRCS -L"$ROOT/gelex_b200" -lggrm_b200 -Wl,-rpath,"$ROOT/gelex_b200"
/tmp/tsan_host 2>&1 | tee /tmp/host_tsan.log | tail -3
echo "TSAN reports: $(grep -c 'WARNING: ThreadSanitizer' /tmp/host_tsan.log)"
rm -f /dev/shm/tsan_w.bin
touch capi.cpp && make -C "$ROOT/gelex_b200/csrc/host" > /dev/null

#include "grm_engine.h"

#include <algorithm>
#include <atomic>
#include <cctype>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <exception>
#include <fstream>
#include <memory>
#include <mutex>
#include <numeric>
#include <thread>
#include <tuple>
#include <unordered_map>
#include <condition_variable>
#include <deque>
#include <map>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/statvfs.h>
#include <sys/vfs.h>
#include <unistd.h>

#include "exceptions.h"
#include "plink_meta.h"

namespace fs = std::filesystem;

namespace gelex::b200 {

namespace {

using Clock = std::chrono::steady_clock;
double since(Clock::time_point t0) { return std::chrono::duration<double>(Clock::now() - t0).count(); }

void check(ggrm_t* h, int status) {
    if (status != GGRM_OK) throw_status(status, ggrm_last_error(h));
}
inline void check_status(ggrm_t* h, int status) { check(h, status); }

struct HandleGuard {
    ggrm_t* h = nullptr;
    ~HandleGuard() { if (h != nullptr) ggrm_destroy(h); }
};

// ---------------------------------------------------------------------------------------------------------------
// .bin writer.  Byte layout = grm_bin_writer.cpp:54-61 (row-major lower triangle incl. diagonal, float32, no header).
// The reference issues one 4-byte ostream::write per element; here rows come off the GPU in blocks (ggrm_fetch /
// ggrm_fetch_block into a ring of page-locked buffers) and a small pool of threads pwrite()s each block at its file
// offset while the next block is already crossing PCIe.  Several GPUs (row slices, or 2D tiles) write disjoint byte
// ranges of the same pre-sized file concurrently.
// Two ways to put a block at its offset (Config::writer / GELEX_B200_WRITER):
//   pwrite  one system call per segment; buffered writes to ONE inode serialise on the inode lock, so several writer threads
//           (and, in the multi-GPU modes, several GPUs' pools) queue behind each other, and the 8 KiB row segments of a 2D-tile
//           run pay a system call each (80 GB at n = 200 000 = 10 M calls)
//   mmap    the pre-sized file is mapped MAP_SHARED and segments are memcpy()d into the mapping: no system call, no inode lock,
//           first-touch page faults run in parallel.  Measured on tmpfs with 4-8 threads: 3.1-3.8 GB/s against 1.3-2.0 GB/s
//           (profiles/r01_summary.md, writer microbenchmark).  A store into a mapped hole of a FULL file system raises SIGBUS
//           instead of returning ENOSPC, so a run only maps its outputs when the file system had room for ALL of them at the
//           start (RunContext::allow_map); otherwise every write goes through pwrite and fails the reference's way.
// Auto = mmap for segments below kMapSegmentBytes (scattered tile rows) and for every segment on tmpfs, pwrite otherwise.
constexpr int64_t kMapSegmentBytes = int64_t{256} << 10;

// Mappings that a Prefaulter has populated, by inode.  A BinFile opened for the same file writes through THAT mapping instead of a
// fresh one of its own: the pages are already in this process's page tables (measured: 23 GB/s, against 5.6 GB/s through a second
// mapping of the same populated file, which still takes a minor fault per page).  A Prefaulter outlives every BinFile of its run.
class Prefaulter;
struct SharedMappings {
    struct Entry { char* map; size_t bytes; Prefaulter* owner; };
    std::mutex mu;
    std::map<std::pair<uint64_t, uint64_t>, Entry> by_inode;
    static SharedMappings& get() {
        static SharedMappings s;
        return s;
    }
    bool find(int fd, size_t bytes, Entry* out) {
        struct stat st {};
        if (::fstat(fd, &st) != 0) return false;
        std::lock_guard<std::mutex> lk(mu);
        const auto it = by_inode.find({static_cast<uint64_t>(st.st_dev), static_cast<uint64_t>(st.st_ino)});
        if (it == by_inode.end() || it->second.bytes != bytes) return false;
        *out = it->second;
        return true;
    }
};
// Blocks until the prefaulter has tried every piece; false = it could not populate the file (no room, old kernel): a store into
// the mapping might then raise SIGBUS, so the caller must write with pwrite, whose ENOSPC becomes FileWriteException.
bool prefault_completed_ok(Prefaulter* p, bool* pinned);

struct BinFile {
    int fd = -1;
    std::string path;
    char* map = nullptr;       // whole file, MAP_SHARED (nullptr: pwrite only)
    size_t map_bytes = 0;
    bool owns_map = true;      // false: the mapping belongs to a Prefaulter (SharedMappings)
    bool map_all = false;      // every segment goes through the mapping (tmpfs, or WriterMode::Mmap)
    bool on_tmpfs = false;
    bool pinned = false;       // the (prepared) mapping is page-locked: ggrm_fetch* may DMA rows straight into it
    BinFile() = default;
    BinFile(const BinFile&) = delete;
    BinFile& operator=(const BinFile&) = delete;
    ~BinFile() {
        if (map != nullptr && owns_map) ::munmap(map, map_bytes);
        if (fd >= 0) ::close(fd);
    }
    // create = truncate + size the file (one caller); otherwise open an existing, already sized file for writing
    // use_prepared: a mapping of this very file that a Prefaulter of this process has fully populated is written through even in
    // Pwrite mode (its pages exist, so no store can fault for lack of space); the caller passes false when the user forced pwrite
    void open(const fs::path& p, int64_t n, bool create, WriterMode mode = WriterMode::Pwrite, bool use_prepared = true) {
        path = p.string();
        const int extra = create ? (O_CREAT | O_TRUNC) : 0;
        if (mode != WriterMode::Pwrite) fd = ::open(path.c_str(), O_RDWR | extra, 0644);  // a shared writable mapping needs read access too
        if (fd < 0) {
            mode = WriterMode::Pwrite;  // e.g. an existing write-only file: the reference only needs to write it, so do we
            fd = ::open(path.c_str(), O_WRONLY | extra, 0644);
        }
        if (fd < 0) throw FileOpenException(path + ": failed to open file");
        const int64_t bytes = n * (n + 1) / 2 * static_cast<int64_t>(sizeof(float));
        if (create && ::ftruncate(fd, static_cast<off_t>(bytes)) != 0)
            throw FileWriteException(path + ": failed to write GRM data to binary file");
        if (bytes == 0) return;
        SharedMappings::Entry shared{};
        if (!create && use_prepared && SharedMappings::get().find(fd, static_cast<size_t>(bytes), &shared)) {
            if (!prefault_completed_ok(shared.owner, &pinned)) return;  // not populated: pwrite only (see prefault_completed_ok)
            map = shared.map;
            owns_map = false;
        } else {
            if (mode == WriterMode::Pwrite) return;
            void* m = ::mmap(nullptr, static_cast<size_t>(bytes), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
            if (m == MAP_FAILED) return;  // e.g. a file system without mmap: pwrite still works
            map = static_cast<char*>(m);
        }
        map_bytes = static_cast<size_t>(bytes);
        struct statfs sf {};
        on_tmpfs = ::fstatfs(fd, &sf) == 0 && static_cast<unsigned long>(sf.f_type) == 0x01021994ul;  // TMPFS_MAGIC
        map_all = mode == WriterMode::Mmap || on_tmpfs || !owns_map;  // a populated mapping takes every segment
    }
    void close_checked() {
        if (map != nullptr) {
            // stores into a mapping of a disk file reach the disk by write-back, whose errors only msync reports (tmpfs has none)
            int rc = on_tmpfs ? 0 : ::msync(map, map_bytes, MS_SYNC);
            if (owns_map && ::munmap(map, map_bytes) != 0) rc = -1;
            map = nullptr;
            if (rc != 0) throw FileWriteException(path + ": failed to write GRM data to binary file");
        }
        if (fd >= 0 && ::close(fd) != 0) {
            fd = -1;
            throw FileWriteException(path + ": failed to write GRM data to binary file");
        }
        fd = -1;
    }
};

// Room for `bytes` more (plus a margin) on the file system that will hold `path`?  Decides once per run whether outputs may be mapped.
bool file_system_has_room(const fs::path& path, int64_t bytes) {
    fs::path dir = fs::absolute(path).parent_path();
    struct statvfs sv {};
    if (::statvfs(dir.c_str(), &sv) != 0) return false;
    const double avail = static_cast<double>(sv.f_bavail) * static_cast<double>(sv.f_frsize);
    return avail >= static_cast<double>(bytes) * 1.02 + static_cast<double>(int64_t{256} << 20);
}

// Page allocation, not copying, is what a fresh output file costs (tmpfs on the build host: 2.5 GB/s into untouched pages, 29 GB/s
// into populated ones; profiles/r01_summary.md §11).  The matrix of a non-LOCO run is only final when the last SNP has been
// accumulated, so the file sits idle for the whole pass: a Prefaulter populates the pages of the pre-sized file meanwhile
// (MADV_POPULATE_WRITE on its own shared mapping, a few threads, 64 MiB at a time; contents untouched, so it may still be running
// when the writers start).  A file system without room makes the call fail instead of raising SIGBUS; the threads then simply stop.
class Prefaulter {
   public:
    Prefaulter(const std::string& path, int threads, bool pin = false) : pin_(pin) {
        fd_ = ::open(path.c_str(), O_RDWR);
        if (fd_ < 0) return;
        struct stat st {};
        if (::fstat(fd_, &st) != 0 || st.st_size <= 0) return;
        bytes_ = static_cast<size_t>(st.st_size);
        void* m = ::mmap(nullptr, bytes_, PROT_READ | PROT_WRITE, MAP_SHARED, fd_, 0);
        if (m == MAP_FAILED) return;
        map_ = static_cast<char*>(m);
        {
            auto& reg = SharedMappings::get();
            std::lock_guard<std::mutex> lk(reg.mu);
            key_ = {static_cast<uint64_t>(st.st_dev), static_cast<uint64_t>(st.st_ino)};
            registered_ = reg.by_inode.emplace(key_, SharedMappings::Entry{map_, bytes_, this}).second;
        }
        populating_.store(std::max(1, threads));
        running_.store(std::max(1, threads));
        for (int t = 0; t < std::max(1, threads); ++t)
            workers_.emplace_back([this] {
                run();
                // the last worker out page-locks the populated mapping (tmpfs only; ~11 GB/s on the GPU box, a disk file's pages
                // cannot be pinned and the call fails): the read-out then DMAs rows straight into the file's pages - no staging
                // buffer, no host copy (measured 56 GB/s against ~10 GB/s through writer threads, profiles/r02_summary.md)
                if (populating_.fetch_sub(1) == 1 && pin_ && !failed_.load() && !stop_.load() && populated_.load() == bytes_)
                    pinned_.store(ggrm_host_register(map_, static_cast<int64_t>(bytes_)) == GGRM_OK);
                running_.fetch_sub(1);
            });
    }
    Prefaulter(const Prefaulter&) = delete;
    Prefaulter& operator=(const Prefaulter&) = delete;
    ~Prefaulter() {
        stop_.store(true);
        for (auto& w : workers_) w.join();
        if (registered_) {
            auto& reg = SharedMappings::get();
            std::lock_guard<std::mutex> lk(reg.mu);
            reg.by_inode.erase(key_);
        }
        if (pinned_.load()) ggrm_host_unregister(map_);
        if (map_ != nullptr) ::munmap(map_, bytes_);
        if (fd_ >= 0) ::close(fd_);
    }
    // until every page has been tried (the page-locking of a pinned prefaulter may still be running: see finished())
    void wait_populated() {
        while (populating_.load() != 0) std::this_thread::sleep_for(std::chrono::microseconds(100));
    }
    bool pinned() const { return pinned_.load(); }
    void wait() {  // (self-test) until every piece has been tried
        for (auto& w : workers_) w.join();
        workers_.clear();
    }
    size_t populated_bytes() const { return populated_.load(); }
    // every piece tried (any thread may ask; does not join) / every piece populated
    bool finished() const { return running_.load() == 0; }
    bool ok() const { return finished() && !failed_.load() && populated_.load() == bytes_; }

   private:
    static constexpr size_t kPiece = size_t{64} << 20;
    void run() {
        for (;;) {
            const size_t off = next_.fetch_add(kPiece);
            if (off >= bytes_ || stop_.load()) return;
            const size_t len = std::min(kPiece, bytes_ - off);
#ifdef MADV_POPULATE_WRITE
            if (::madvise(map_ + off, len, MADV_POPULATE_WRITE) != 0) {  // old kernel, or no room: the writers fall back to pwrite
                failed_.store(true);
                return;
            }
            populated_.fetch_add(len);
#else
            failed_.store(true);
            return;
#endif
        }
    }
    int fd_ = -1;
    char* map_ = nullptr;
    size_t bytes_ = 0;
    std::atomic<size_t> next_{0}, populated_{0};
    std::atomic<bool> stop_{false}, failed_{false}, pinned_{false};
    std::atomic<int> running_{0}, populating_{0};
    bool pin_ = false;
    std::vector<std::thread> workers_;
    std::pair<uint64_t, uint64_t> key_{0, 0};
    bool registered_ = false;
};

bool prefault_completed_ok(Prefaulter* p, bool* pinned) {
    if (p == nullptr) return false;
    while (!p->finished()) std::this_thread::sleep_for(std::chrono::microseconds(100));
    if (pinned != nullptr) *pinned = p->pinned();
    return p->ok();
}

struct Segment { int64_t src_elem, file_byte, bytes; };  // buffer elements [src_elem, +bytes/4) -> file bytes [file_byte, +bytes)

class WriterPool;
// A pool that comes into being at its first use: its page-locked buffers (6 x 16 MB; ~0.5 ms per MB, serialised by the driver across
// the GPU threads of one process) are then allocated while the GPUs accumulate, not in front of the first push.
class LazyPool {
   public:
    LazyPool(int threads, int buffers, int64_t buffer_elems) : threads_(threads), buffers_(buffers), elems_(buffer_elems) {}
    WriterPool& get();
    WriterPool* peek() { return pool_.get(); }

   private:
    int threads_, buffers_;
    int64_t elems_;
    std::unique_ptr<WriterPool> pool_;
};

class WriterPool {
   public:
    WriterPool(int threads, int buffers, int64_t buffer_elems) : elems_(buffer_elems) {
        for (int i = 0; i < buffers; ++i) {
            void* p = nullptr;
            if (ggrm_host_alloc(buffer_elems * static_cast<int64_t>(sizeof(float)), &p) != GGRM_OK || p == nullptr) {
                pageable_.emplace_back(static_cast<size_t>(buffer_elems));  // pinned memory exhausted: slower, still correct
                p = pageable_.back().data();
            } else {
                pinned_.push_back(p);
            }
            free_.push_back(static_cast<float*>(p));
        }
        for (int t = 0; t < std::max(1, threads); ++t) workers_.emplace_back([this] { run(); });
    }
    ~WriterPool() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_job_.notify_all();
        for (auto& w : workers_) w.join();
        for (void* p : pinned_) ggrm_host_free(p);
    }
    int64_t buffer_elems() const { return elems_; }
    float* acquire() {
        std::unique_lock<std::mutex> lk(mu_);
        cv_free_.wait(lk, [this] { return !free_.empty() || failed_; });
        if (failed_) throw FileWriteException(failed_path_ + ": failed to write GRM data to binary file");
        float* b = free_.back();
        free_.pop_back();
        return b;
    }
    void submit(const BinFile& file, float* buf, std::vector<Segment> segs) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            jobs_.push_back({file.fd, &file.path, buf, std::move(segs), file.map, file.map_all});
            ++inflight_;
        }
        cv_job_.notify_one();
    }
    void drain() {
        std::unique_lock<std::mutex> lk(mu_);
        cv_free_.wait(lk, [this] { return inflight_ == 0; });
        if (failed_) throw FileWriteException(failed_path_ + ": failed to write GRM data to binary file");
    }
    double busy_s() const { return busy_s_; }

   private:
    struct Job { int fd; const std::string* path; float* buf; std::vector<Segment> segs; char* map; bool map_all; };
    void run() {
        for (;;) {
            Job job;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_job_.wait(lk, [this] { return stop_ || !jobs_.empty(); });
                if (jobs_.empty()) return;
                job = std::move(jobs_.front());
                jobs_.pop_front();
            }
            const auto t0 = Clock::now();
            bool ok = true;
            for (const Segment& sg : job.segs) {
                const char* src = reinterpret_cast<const char*>(job.buf + sg.src_elem);
                if (job.map != nullptr && (job.map_all || sg.bytes < kMapSegmentBytes)) {
                    std::memcpy(job.map + sg.file_byte, src, static_cast<size_t>(sg.bytes));
                    continue;
                }
                int64_t done = 0;
                while (ok && done < sg.bytes) {
                    const ssize_t w = ::pwrite(job.fd, src + done, static_cast<size_t>(sg.bytes - done), static_cast<off_t>(sg.file_byte + done));
                    if (w <= 0) ok = false; else done += w;
                }
                if (!ok) break;
            }
            const double dt = since(t0);
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (!ok && !failed_) {
                    failed_ = true;
                    failed_path_ = *job.path;
                }
                busy_s_ += dt;
                free_.push_back(job.buf);
                --inflight_;
            }
            cv_free_.notify_all();
        }
    }
    std::string failed_path_;
    int64_t elems_;
    std::mutex mu_;
    std::condition_variable cv_job_, cv_free_;
    std::deque<Job> jobs_;
    std::vector<float*> free_;
    std::vector<void*> pinned_;
    std::vector<std::vector<float>> pageable_;
    std::vector<std::thread> workers_;
    int inflight_ = 0;
    bool stop_ = false, failed_ = false;
    double busy_s_ = 0;
};

WriterPool& LazyPool::get() {
    if (!pool_) pool_ = std::make_unique<WriterPool>(threads_, buffers_, elems_);
    return *pool_;
}

inline int64_t tri(int64_t i) { return i * (i + 1) / 2; }

// Largest count of whole rows starting at r (not past r1) whose packed entries fit `cap` elements; a single row always fits
// (buffers hold >= n floats).
int64_t rows_fitting(int64_t r, int64_t r1, int64_t cap) {
    int64_t lo = 1, hi = r1 - r;
    while (lo < hi) {
        const int64_t mid = (lo + hi + 1) / 2;
        if (tri(r + mid) - tri(r) <= cap) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// Row segments of the rows x cols block at (row0, col0), held row-major with pitch `cols`, that lie in the lower triangle.
std::vector<Segment> supertile_segments(int64_t row0, int64_t col0, int64_t rows, int64_t cols) {
    std::vector<Segment> segs;
    segs.reserve(static_cast<size_t>(rows));
    for (int64_t i = 0; i < rows; ++i) {
        const int64_t gi = row0 + i;
        const int64_t len = std::min(cols, gi + 1 - col0);
        if (len > 0) segs.push_back(Segment{i * cols, (tri(gi) + col0) * 4, len * 4});
    }
    return segs;
}

// Rows [r0, r1) of matrix (kind, group) -> their byte range of the file, in blocks of whole rows.
void write_packed_rows(ggrm_t* h, int kind, int group, int64_t r0, int64_t r1, const BinFile& file, WriterPool& pool, bool gcta = false) {
    int64_t r = r0;
    if (file.map != nullptr && file.pinned) {
        // the file's pages are populated and page-locked: the device copies the packed rows to their place in the file
        constexpr int64_t kDirectElems = int64_t{1} << 28;  // 1 GiB per call keeps the progress of a failing run observable
        while (r < r1) {
            const int64_t cnt = rows_fitting(r, r1, kDirectElems);
            float* dst = reinterpret_cast<float*>(file.map + tri(r) * 4);
            const int st = gcta ? ggrm_fetch_gcta(h, kind, group, r, cnt, dst) : ggrm_fetch(h, kind, group, r, cnt, dst, nullptr);
            if (st != GGRM_OK) throw_status(st, ggrm_last_error(h));
            r += cnt;
        }
        return;
    }
    while (r < r1) {
        const int64_t cnt = rows_fitting(r, r1, pool.buffer_elems());
        float* buf = pool.acquire();
        // gcta: kind 0 -> A = G / N, kind 1 -> N (ggrm_fetch_gcta's `which` is the kind index)
        const int st = gcta ? ggrm_fetch_gcta(h, kind, group, r, cnt, buf) : ggrm_fetch(h, kind, group, r, cnt, buf, nullptr);
        if (st != GGRM_OK) {
            pool.submit(file, buf, {});
            throw_status(st, ggrm_last_error(h));
        }
        pool.submit(file, buf, {Segment{0, tri(r) * 4, (tri(r + cnt) - tri(r)) * 4}});
        r += cnt;
    }
}

// 2D-tile mode: every 2048 x 2048 super-tile this handle owns -> row segments at their own offsets.
void write_owned_supertiles(ggrm_t* h, int kind, int group, int64_t n, const BinFile& file, WriterPool& pool, bool gcta = false) {
    constexpr int64_t kSuper = 2048;
    int64_t count = 0;
    check_status(h, ggrm_owned_supertiles(h, nullptr, nullptr, 0, &count));
    std::vector<int64_t> sr(static_cast<size_t>(std::max<int64_t>(1, count))), sc(sr.size());
    check_status(h, ggrm_owned_supertiles(h, sr.data(), sc.data(), count, &count));
    for (int64_t t = 0; t < count; ++t) {
        const int64_t row0 = sr[static_cast<size_t>(t)] * kSuper, col0 = sc[static_cast<size_t>(t)] * kSuper;
        const int64_t rows = std::min(kSuper, n - row0), cols = std::min(kSuper, n - col0);
        if (rows <= 0 || cols <= 0) continue;
        float* buf = pool.acquire();
        const int st = gcta ? ggrm_fetch_block_gcta(h, kind, group, row0, rows, col0, cols, buf, cols) : ggrm_fetch_block(h, kind, group, row0, rows, col0, cols, buf, cols);
        if (st != GGRM_OK) {
            pool.submit(file, buf, {});
            throw_status(st, ggrm_last_error(h));
        }
        pool.submit(file, buf, supertile_segments(row0, col0, rows, cols));
    }
}

// SNP shards of a GCTA run with the default diagonal: total diagonal sums of handle r = its own + every peer's (the peers'
// passes must be complete - a band flag of theirs is up, or they were finished).
void reduce_gcta_diag(const std::vector<ggrm_t*>& handles, size_t r) {
    std::vector<const void*> peers;
    for (size_t p = 0; p < handles.size(); ++p) {
        if (p == r) continue;
        void* ptr = nullptr;
        check(handles[p], ggrm_gcta_diag_address(handles[p], &ptr));
        peers.push_back(ptr);
    }
    if (peers.size() > 7) throw GelexException("GCTA diagonal: more than 8 GPUs in one SNP-shard run");
    check(handles[r], ggrm_gcta_diag_reduce_async(handles[r], peers.data(), static_cast<int>(peers.size())));
}

std::string lower(std::string s) {
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return static_cast<char>(std::tolower(c)); });
    return s;
}

}  // namespace

GenotypeProcessMethod parse_genotype_process_method(const std::string& value) {
    static const std::unordered_map<std::string, GenotypeProcessMethod> kMap = {
        {"standardizehwe", GenotypeProcessMethod::StandardizeHWE}, {"sh", GenotypeProcessMethod::StandardizeHWE},
        {"centerhwe", GenotypeProcessMethod::CenterHWE}, {"ch", GenotypeProcessMethod::CenterHWE},
        {"orthstandardizehwe", GenotypeProcessMethod::OrthStandardizeHWE}, {"osh", GenotypeProcessMethod::OrthStandardizeHWE},
        {"orthcenterhwe", GenotypeProcessMethod::OrthCenterHWE}, {"och", GenotypeProcessMethod::OrthCenterHWE},
        {"standardize", GenotypeProcessMethod::Standardize}, {"s", GenotypeProcessMethod::Standardize},
        {"center", GenotypeProcessMethod::Center}, {"c", GenotypeProcessMethod::Center},
        {"orthstandardize", GenotypeProcessMethod::OrthStandardize}, {"os", GenotypeProcessMethod::OrthStandardize},
        {"orthcenter", GenotypeProcessMethod::OrthCenter}, {"oc", GenotypeProcessMethod::OrthCenter},
    };
    const auto it = kMap.find(lower(value));
    if (it == kMap.end())
        throw InvalidInputException("Invalid genotype process method: \"" + value +
                                    "\". Valid: StandardizeHWE(SH), CenterHWE(CH), OrthStandardizeHWE(OSH), OrthCenterHWE(OCH),"
                                    " Standardize(S), Center(C), OrthStandardize(OS), OrthCenter(OC)");
    return it->second;
}

const char* method_name(GenotypeProcessMethod m) {
    switch (m) {
        case GenotypeProcessMethod::StandardizeHWE: return "StandardizeHWE";
        case GenotypeProcessMethod::CenterHWE: return "CenterHWE";
        case GenotypeProcessMethod::OrthStandardizeHWE: return "OrthStandardizeHWE";
        case GenotypeProcessMethod::OrthCenterHWE: return "OrthCenterHWE";
        case GenotypeProcessMethod::Standardize: return "Standardize";
        case GenotypeProcessMethod::Center: return "Center";
        case GenotypeProcessMethod::OrthStandardize: return "OrthStandardize";
        case GenotypeProcessMethod::OrthCenter: return "OrthCenter";
    }
    return "Unknown Genotype Process Method";
}

void write_id_file(const fs::path& path, const std::vector<std::string>& sorted_ids) {
    std::ofstream f(path, std::ios::out | std::ios::trunc);
    if (!f.is_open()) throw FileOpenException(path.string() + ": failed to open file");
    for (const auto& id : sorted_ids) {
        const auto [fid, iid] = split_sample_id(id);
        f << fid << '\t' << iid << '\n';
    }
    f.flush();
    if (!f.good()) throw FileWriteException(path.string() + ": failed to write ID data");
}

std::vector<GrmWorkItem> GrmEngine::build_plan(const std::vector<std::string>& chroms, GrmType mode, bool do_loco,
                                               std::string* output_pattern_suffix, size_t* num_groups) {
    std::vector<std::pair<std::string, bool>> tasks;  // build_tasks, grm_work_plan.cpp:38-50
    if (mode != GrmType::D) tasks.emplace_back("add", true);
    if (mode != GrmType::A) tasks.emplace_back("dom", false);
    const std::string task_pattern = tasks.size() == 1 ? tasks[0].first : std::string("{add|dom}");
    std::vector<GrmWorkItem> items;
    const int64_t m = static_cast<int64_t>(chroms.size());
    if (!do_loco) {
        for (const auto& t : tasks) items.push_back({{{0, m}}, t.second, t.first, -1});
        if (num_groups != nullptr) *num_groups = 0;
        if (output_pattern_suffix != nullptr) *output_pattern_suffix = "." + task_pattern + ".{bin|id}";
    } else {
        const auto runs = chromosome_runs(chroms);
        for (size_t g = 0; g < runs.size(); ++g)
            for (const auto& t : tasks)
                items.push_back({{{runs[g].begin, runs[g].end}}, t.second, t.first + ".chr" + runs[g].name, static_cast<int>(g)});
        if (num_groups != nullptr) *num_groups = runs.size();
        if (output_pattern_suffix != nullptr)
            *output_pattern_suffix = "." + task_pattern + ".chr{1.." + std::to_string(runs.size()) + "}.{bin|id}";
    }
    return items;
}

GrmEngine::GrmEngine(Config config) : config_(std::move(config)) {}

namespace {

constexpr int64_t kSuperRows = 2048;          // ownership / row-slice granularity of the device library (8 tiles of 256)
constexpr int64_t kWriterBufferElems = int64_t{4} << 20;  // 16 MiB: one 2048 x 2048 super-tile, or ~80 rows at n = 50 000

struct RunContext {
    const GrmEngine::Config* cfg;
    const SampleOrder* order;
    const BedMapping* bed;
    int64_t n_raw, n_out, m, bpv;
    std::vector<ChrRun> groups;  // one pseudo-group [0, m) when not LOCO
    bool loco;
    bool want_add, want_dom;
    bool gcta;  // kind 1 is the non-missing pair count matrix; outputs are <out>.grm.bin / .grm.N.bin / .grm.id
    bool gcta_diag = false;  // ... and the diagonal of .grm.bin is GCTA's default one (GGRM_SECOND_NONMISSING_GCTA_DIAG)
    int precision;  // operand precision of this run (Config::precision unless FP16 cannot represent the code values, see compute())
    WriterMode writer;  // effective mode of this run: Pwrite when the file system could not hold every output (see BinFile)
    bool bed_relay;     // 2D-tile mode: one upload per host chunk, peers copy it over NVLink (BedRelay)
    bool prefault;      // create output files ahead of the writers and populate their pages meanwhile (Prefaulter / OutputPreparer);
                        // false when the user forced pwrite.  Independent of `writer`: a populated mapping cannot fault for lack of space
    bool direct_dma;    // prepared mappings are page-locked and whole-row read-outs DMA straight into them (GELEX_B200_DIRECT_DMA=1; off by default)
    // progress (reference: per work item, counted in SNPs; grm.h:124-130)
    const GrmObserver* observer;
    std::mutex* mu;
    std::vector<std::string>* files;
    double* write_s;
    Clock::time_point t_start;
    std::vector<GrmEngine::Summary::Phase>* phases;
};

// Record a host-side phase (seconds since compute() began); per name the earliest begin and the latest end over the GPU threads.
void note_phase(const RunContext& rc, const char* name, Clock::time_point a, Clock::time_point b) {
    const double ta = std::chrono::duration<double>(a - rc.t_start).count(), tb = std::chrono::duration<double>(b - rc.t_start).count();
    std::lock_guard<std::mutex> lk(*rc.mu);
    for (auto& p : *rc.phases)
        if (p.name == name) {
            p.begin = std::min(p.begin, ta);
            p.end = std::max(p.end, tb);
            return;
        }
    rc.phases->push_back({name, ta, tb});
}

// Which part of an output file a handle contributes.
struct WritePart {
    enum Kind { AllRows, RowSlice, OwnedTiles } kind = AllRows;
    int64_t r0 = 0, r1 = 0;   // RowSlice
    bool create = true;        // this caller creates + sizes the .bin and writes the .id (otherwise the main thread did)
    bool precreated = false;   // with create: the .bin already exists, sized (and populated) by an OutputPreparer; .id / registration still ours
};

ggrm_t* create_handle(const RunContext& rc, int device, const std::vector<int>& my_groups, bool keep_whole, int tile_rank = 0, int tile_world = 1) {
    ggrm_config_t c{};
    c.abi_version = GGRM_ABI_VERSION;
    c.device = device;
    c.n_raw = rc.n_raw;
    c.n_out = rc.n_out;
    c.raw_to_out = rc.order->dense ? nullptr : rc.order->raw_to_out.data();
    c.m_total = rc.m;
    c.method = static_cast<int>(rc.cfg->method);
    c.want_add = rc.want_add;
    c.want_dom = rc.want_dom;
    c.precision = rc.precision;
    std::vector<int64_t> gb, ge;
    for (int g : my_groups) {
        gb.push_back(rc.groups[static_cast<size_t>(g)].begin);
        ge.push_back(rc.groups[static_cast<size_t>(g)].end);
    }
    c.n_groups = static_cast<int>(my_groups.size());
    c.group_begin = gb.data();
    c.group_end = ge.data();
    c.keep_whole = keep_whole ? 1 : 0;
    c.all_missing_policy = rc.cfg->skip_all_missing ? GGRM_ALLMISS_SKIP : GGRM_ALLMISS_REFERENCE_NAN;
    c.tile_rank = tile_rank;
    c.tile_world = tile_world;
    c.second_matrix = rc.gcta ? (rc.gcta_diag ? GGRM_SECOND_NONMISSING_GCTA_DIAG : GGRM_SECOND_NONMISSING) : GGRM_SECOND_DOMINANCE;
    ggrm_t* h = nullptr;
    const int st = ggrm_create(&c, &h);
    if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
    return h;
}

// Progress of a work item.  The reference runs the dominance item of a SNP range AFTER its additive item (grm_work_plan.cpp:94-103,
// 139-149: per group, one item per task) and each item reports (current, total) chunk by chunk (grm.h:124-130).  One pass computes
// both matrices here, so the additive item's events go out as the pass advances and the dominance item's - the same sequence once
// more - are replayed when the pass over the range is complete: an observer sees the events of the reference in the reference's order.
using ProgressReplay = std::vector<std::pair<size_t, size_t>>;

void emit_progress_locked(const RunContext& rc, size_t current, size_t total, ProgressReplay& replay) {
    (*rc.observer)(GrmProgressEvent{current, total, false});
    const int kinds = (rc.want_add ? 1 : 0) + (rc.want_dom && !rc.gcta ? 1 : 0);
    if (kinds < 2) return;
    if (!replay.empty() && current <= replay.back().first) replay.clear();  // an item that was abandoned (exception) left its events behind
    replay.emplace_back(current, total);
    if (current == total) {
        for (const auto& e : replay) (*rc.observer)(GrmProgressEvent{e.first, e.second, false});
        replay.clear();
    }
}

void emit_progress(const RunContext& rc, size_t current, size_t total) {
    if (rc.observer == nullptr || !*rc.observer) return;
    thread_local ProgressReplay replay;  // a GPU thread works through its items one after the other
    std::lock_guard<std::mutex> lk(*rc.mu);
    emit_progress_locked(rc, current, total, replay);
}

// SNP shards of ONE work item on several GPUs: the observer was told the item has `total` SNPs (GrmComputeStartedEvent), so the
// ranks report against one shared counter - (sum over ranks, total), monotone because the counter moves under the observer's lock.
void emit_progress_shared(const RunContext& rc, int64_t* shared_done, int64_t step, size_t total, ProgressReplay* replay) {
    std::lock_guard<std::mutex> lk(*rc.mu);
    *shared_done += step;
    if (rc.observer == nullptr || !*rc.observer) return;
    ProgressReplay none;
    emit_progress_locked(rc, static_cast<size_t>(*shared_done), total, replay != nullptr ? *replay : none);
}

// ---------------------------------------------------------------------------------------------------------------
// One host read of the BED for every GPU of a 2D-tile run (Config::bed_relay; SURVEY 8e, input side of the tile partition).
// All GPUs need all SNPs.  Without the relay each GPU thread pushes the whole mapped file through its own pinned ring and PCIe
// link (C5: 30 GB per GPU, 5.1 s of accumulate for 1.85 s of SYRK, profiles/r01_summary.md §5).  With it, host chunk j is uploaded
// ONCE - by an uploader thread of rank j % N into one of that GPU's staging slots (ggrm_stage_upload) - and every rank takes it
// from there with ggrm_push_device: a copy over NVLink peer memory for the N - 1 others.  All ranks walk the same chunk list in
// the same order (run_groups_on_device cuts every chromosome run from its first SNP in steps of chunk_size), so a chunk is its
// index; slot (j / N) % slots of rank j % N is reused once all N ranks have consumed chunk j - N * slots.  No collective: flags
// in process memory, as in the band-pipelined reduce.
// The protocol, free of any device call (GrmEngine::relay_selftest drives it with host buffers): `world` uploader threads fill
// slots, `world` consumers take every chunk in order.
class ChunkRelay {
   public:
    using Upload = std::function<void(size_t owner, int slot, size_t chunk)>;  // blocking: on return the slot holds the chunk
    ChunkRelay(size_t n_chunks, size_t world, size_t slots, Upload upload)
        : n_(n_chunks), world_(world), slots_(slots), upload_(std::move(upload)), ready_(n_chunks), consumed_(n_chunks), errors_(world) {
        for (size_t r = 0; r < world_; ++r) uploaders_.emplace_back([this, r] { upload_loop(r); });
    }
    ChunkRelay(const ChunkRelay&) = delete;
    ChunkRelay& operator=(const ChunkRelay&) = delete;
    ~ChunkRelay() { stop(); }
    size_t owner_of(size_t j) const { return j % world_; }
    int slot_of(size_t j) const { return static_cast<int>((j / world_) % slots_); }
    // Consumer side: block until chunk j sits in slot_of(j) of owner_of(j) ...
    void acquire(size_t j) {
        wait_for([&] { return ready_[j].load(std::memory_order_acquire) != 0; }, "BED relay: timed out waiting for a chunk");
    }
    // ... and, once this consumer has copied it out, let the slot go (it is refilled when all `world` consumers have).
    void release(size_t j) { consumed_[j].fetch_add(1, std::memory_order_release); }
    void fail() { failed_.store(true); }
    // Joins the uploaders (after the consumers have finished or failed) and rethrows the first uploader error, if any.
    void stop_and_rethrow() {
        stop();
        for (auto& e : errors_) if (e) std::rethrow_exception(e);
    }

   private:
    void stop() {
        failed_.store(true);  // a no-op after a complete run: every uploader has left its loop
        for (auto& t : uploaders_) t.join();
        uploaders_.clear();
    }
    template <typename Pred>
    void wait_for(Pred ok, const char* what) {
        const auto t0 = Clock::now();
        while (!ok()) {
            if (failed_.load()) throw GelexException("multi-GPU run aborted: another GPU failed");
            if (since(t0) > 3600.0) throw GelexException(what);
            std::this_thread::sleep_for(std::chrono::microseconds(20));
        }
    }
    void upload_loop(size_t r) {
        try {
            const size_t lap = world_ * slots_;
            for (size_t j = r; j < n_; j += world_) {
                if (j >= lap)  // the chunk that last used this slot has been taken by every consumer
                    wait_for([&] { return consumed_[j - lap].load(std::memory_order_acquire) == static_cast<int>(world_); },
                             "BED relay: timed out waiting for a free slot");
                upload_(r, slot_of(j), j);
                ready_[j].store(1, std::memory_order_release);
            }
        } catch (...) {
            if (!failed_.exchange(true)) errors_[r] = std::current_exception();  // only a first failure is a cause
        }
    }
    size_t n_, world_, slots_;
    Upload upload_;
    std::vector<std::atomic<int>> ready_, consumed_;
    std::vector<std::exception_ptr> errors_;
    std::vector<std::thread> uploaders_;
    std::atomic<bool> failed_{false};
};

class BedRelay {
   public:
    BedRelay(const RunContext& rc, const std::vector<int>& gpus, int slots_per_gpu) : rc_(rc) {
        for (const ChrRun& g : rc.groups)
            for (int64_t s = g.begin; s < g.end; s += rc.cfg->chunk_size) {
                index_of_[s] = chunks_.size();
                chunks_.push_back({s, std::min<int64_t>(rc.cfg->chunk_size, g.end - s)});
            }
        const size_t world = gpus.size();
        for (size_t a = 0; a < world; ++a)
            for (size_t b = 0; b < world; ++b)
                if (gpus[a] != gpus[b]) {
                    const int st = ggrm_enable_peer_access(gpus[a], gpus[b]);
                    if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
                }
        stages_.assign(world, nullptr);
        try {
            for (size_t r = 0; r < world; ++r) {
                const int st = ggrm_stage_create(gpus[r], std::max<int64_t>(1, static_cast<int64_t>(rc.cfg->chunk_size) * rc.bpv), slots_per_gpu, &stages_[r]);
                if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
            }
            relay_ = std::make_unique<ChunkRelay>(chunks_.size(), world, static_cast<size_t>(slots_per_gpu), [this](size_t owner, int slot, size_t j) {
                const Chunk& ck = chunks_[j];
                const uint8_t* p = rc_.bed->chunk_ptr(ck.s, ck.c);
                if (p == nullptr) throw FileFormatException("BedPipe::load_chunk: mapped BED payload is truncated");
                const int st = ggrm_stage_upload(stages_[owner], slot, p, ck.c * rc_.bpv);
                if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
            });
        } catch (...) {
            for (ggrm_stage_t* st : stages_) if (st != nullptr) ggrm_stage_destroy(st);
            throw;
        }
    }
    BedRelay(const BedRelay&) = delete;
    BedRelay& operator=(const BedRelay&) = delete;
    ~BedRelay() {
        relay_.reset();  // joins the uploaders before their slots go away
        for (ggrm_stage_t* st : stages_) if (st != nullptr) ggrm_stage_destroy(st);
    }
    void fail() { relay_->fail(); }
    void stop_and_rethrow() { relay_->stop_and_rethrow(); }
    // The handle of one rank takes the host chunk that starts at SNP s (one of the relay's chunks, c SNPs).
    void push(ggrm_t* h, int64_t s, int64_t c) {
        const auto it = index_of_.find(s);
        if (it == index_of_.end() || chunks_[it->second].c != c) throw GelexException("BED relay: chunk list out of step");
        const size_t j = it->second;
        relay_->acquire(j);
        const uint8_t* d = nullptr;
        const int st = ggrm_stage_slot(stages_[relay_->owner_of(j)], relay_->slot_of(j), &d);
        if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
        check(h, ggrm_push_device(h, d, s, c, rc_.bpv));
        check(h, ggrm_wait_copies(h));  // the slot has been read by this rank
        relay_->release(j);
    }

   private:
    struct Chunk { int64_t s, c; };
    const RunContext& rc_;
    std::vector<Chunk> chunks_;
    std::unordered_map<int64_t, size_t> index_of_;
    std::vector<ggrm_stage_t*> stages_;
    std::unique_ptr<ChunkRelay> relay_;
};

// Push SNPs [begin, end) in host chunks of cfg.chunk_size (the reference's load_chunk granularity); through the relay when there is one.
void push_range(const RunContext& rc, ggrm_t* h, int64_t begin, int64_t end, int64_t item_total, int64_t* item_done, bool report = true,
                BedRelay* relay = nullptr, int64_t* shared_done = nullptr, ProgressReplay* shared_replay = nullptr) {
    for (int64_t s = begin; s < end; s += rc.cfg->chunk_size) {
        const int64_t c = std::min<int64_t>(rc.cfg->chunk_size, end - s);
        if (s < 0 || s + c > rc.m || c <= 0)  // validate_chunk_range, bed_pipe.cpp:38-52
            throw ColumnRangeException("invalid chunk range: [" + std::to_string(s) + ", " + std::to_string(s + c) + "). Total SNPs: " + std::to_string(rc.m));
        if (relay != nullptr) {
            relay->push(h, s, c);
        } else {
            const uint8_t* p = rc.bed->chunk_ptr(s, c);
            if (p == nullptr) throw FileFormatException("BedPipe::load_chunk: mapped BED payload is truncated");
            check(h, ggrm_push(h, p, s, c, rc.bpv));
        }
        *item_done += c;
        if (report && shared_done != nullptr) emit_progress_shared(rc, shared_done, c, static_cast<size_t>(item_total), shared_replay);
        else if (report) emit_progress(rc, static_cast<size_t>(*item_done), static_cast<size_t>(item_total));
    }
}

std::string output_base(const RunContext& rc, int kind, const std::string& suffix) {
    if (rc.gcta) return rc.cfg->out_prefix + (kind == 0 ? ".grm" : ".grm.N");
    return rc.cfg->out_prefix + "." + (kind == 0 ? "add" : "dom") + suffix;
}

bool has_id_file(const RunContext& rc, int kind) { return !(rc.gcta && kind == 1); }  // GCTA: one .grm.id beside both .bin files

void register_files(const RunContext& rc, const std::string& base, bool with_id = true) {
    std::lock_guard<std::mutex> lk(*rc.mu);
    rc.files->push_back(base + ".bin");
    if (with_id) rc.files->push_back(base + ".id");
}

// Main-thread step of the multi-writer modes: size every .bin, write every .id, before any GPU thread opens them.
void precreate_outputs(const RunContext& rc, const std::vector<std::string>& suffixes) {
    for (const std::string& suffix : suffixes)
        for (int kind = 0; kind < 2; ++kind) {
            if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
            const std::string base = output_base(rc, kind, suffix);
            BinFile f;
            f.open(base + ".bin", rc.n_out, true);
            f.close_checked();
            if (has_id_file(rc, kind)) write_id_file(base + ".id", rc.order->ids);
            register_files(rc, base, has_id_file(rc, kind));
        }
}

// Prefaulters for the whole-genome .bin of every requested kind (files must exist, pre-sized).  Empty unless the run maps its
// outputs, asked for it, and is not a LOCO run (whose files are written one after another while the GPU is already ahead).
std::vector<std::unique_ptr<Prefaulter>> start_prefault(const RunContext& rc, bool whole_of_loco_run = false) {
    std::vector<std::unique_ptr<Prefaulter>> out;
    if (!rc.prefault || (rc.loco && !whole_of_loco_run)) return out;
    for (int kind = 0; kind < 2; ++kind)
        if (kind == 0 ? rc.want_add : rc.want_dom)
            out.push_back(std::make_unique<Prefaulter>(output_base(rc, kind, "") + ".bin", std::max(2, rc.cfg->writer_threads), rc.direct_dma));
    return out;
}

// LOCO runs write one 5 GB file per chromosome, one after another, and page allocation - not copying - is what a fresh file
// costs (see Prefaulter).  An OutputPreparer walks the run's output list a few files ahead of the writer: it creates and sizes
// file i + 1 .. i + ahead and populates their pages on its own threads while the writers copy file i into pages that are
// already there.  The populated mapping stays registered (SharedMappings) until the writer reports the file done, so the
// BinFile of that file writes through it.  A preparation that fails (no room, old kernel) leaves the file to the writer, which
// then creates it itself and fails the reference's way.
class OutputPreparer {
   public:
    OutputPreparer(std::vector<std::vector<std::string>> bins /* per entry: the .bin paths written together */, int64_t n, int threads, int ahead, bool pin)
        : bins_(std::move(bins)), n_(n), threads_(std::max(1, threads)), ahead_(std::max(1, ahead)), pin_(pin), ready_(bins_.size()), ok_(bins_.size()), slots_(bins_.size()) {
        for (auto& r : ready_) r.store(0);
        worker_ = std::thread([this] { run(); });
    }
    OutputPreparer(const OutputPreparer&) = delete;
    OutputPreparer& operator=(const OutputPreparer&) = delete;
    ~OutputPreparer() {
        stop_.store(true);
        if (worker_.joinable()) worker_.join();
    }
    // Blocks until entry i has been prepared (or given up on); true = the files exist, sized, pages populated.
    bool wait_ready(size_t i) {
        while (ready_[i].load(std::memory_order_acquire) == 0) std::this_thread::sleep_for(std::chrono::microseconds(50));
        return ok_[i] != 0;
    }
    // The writer has closed entry i's files: the preparer may move on; it also unmaps the populated mappings of finished entries
    // on its own thread (tearing down the page tables of a 5 GB mapping takes ~0.1 s that the pushing thread should not pay).
    void done(size_t i) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            finished_.push_back(i);
        }
        done_.fetch_add(1);
    }

   private:
    void release_finished() {
        std::vector<std::vector<std::unique_ptr<Prefaulter>>> drop;
        {
            std::lock_guard<std::mutex> lk(mu_);
            for (size_t i : finished_) drop.push_back(std::move(slots_[i]));
            finished_.clear();
        }
    }
    void run() {
        for (size_t i = 0; i < bins_.size() && !stop_.load(); ++i) {
            while (i >= done_.load() + static_cast<size_t>(ahead_) && !stop_.load()) {
                release_finished();
                std::this_thread::sleep_for(std::chrono::microseconds(200));
            }
            release_finished();
            if (stop_.load()) break;
            bool ok = true;
            std::vector<std::unique_ptr<Prefaulter>> made;
            try {
                for (const std::string& path : bins_[i]) {
                    BinFile f;
                    f.open(path, n_, true);
                    f.close_checked();
                    made.push_back(std::make_unique<Prefaulter>(path, threads_, pin_));
                    made.back()->wait_populated();  // its page-locking (pin_) runs on while the next file is populated
                }
            } catch (...) {
                ok = false;  // the writer creates the file itself and reports the error the reference would
                made.clear();
            }
            {
                std::lock_guard<std::mutex> lk(mu_);
                slots_[i] = std::move(made);
            }
            ok_[i] = ok ? 1 : 0;
            ready_[i].store(1, std::memory_order_release);
        }
        for (auto& r : ready_) {  // stopped early: nobody may wait forever
            int z = 0;
            r.compare_exchange_strong(z, 1);
        }
        while (!stop_.load()) {  // the last entries are finished after the last one was prepared
            release_finished();
            std::this_thread::sleep_for(std::chrono::microseconds(500));
        }
        release_finished();
    }
    std::vector<std::vector<std::string>> bins_;
    int64_t n_;
    int threads_, ahead_;
    bool pin_;
    std::vector<std::atomic<int>> ready_;
    std::vector<char> ok_;
    std::vector<std::vector<std::unique_ptr<Prefaulter>>> slots_;
    std::vector<size_t> finished_;
    std::atomic<size_t> done_{0};
    std::atomic<bool> stop_{false};
    std::mutex mu_;
    std::thread worker_;
};

void write_outputs(const RunContext& rc, ggrm_t* h, int local_group, const std::string& suffix /* "" or ".chrNAME" */, const WritePart& part,
                   WriterPool& pool) {
    const auto t0 = Clock::now();
    for (int kind = 0; kind < 2; ++kind) {
        if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
        const std::string base = output_base(rc, kind, suffix);
        BinFile f;
        f.open(base + ".bin", rc.n_out, part.create && !part.precreated, rc.writer, rc.prefault);
        try {
            switch (part.kind) {
                case WritePart::AllRows: write_packed_rows(h, kind, local_group, 0, rc.n_out, f, pool, rc.gcta); break;
                case WritePart::RowSlice: write_packed_rows(h, kind, local_group, part.r0, part.r1, f, pool, rc.gcta); break;
                case WritePart::OwnedTiles: write_owned_supertiles(h, kind, local_group, rc.n_out, f, pool, rc.gcta); break;
            }
        } catch (...) {
            try { pool.drain(); } catch (...) {}  // no writer may still hold the descriptor when it closes
            throw;
        }
        pool.drain();
        f.close_checked();
        if (part.create) {
            if (has_id_file(rc, kind)) write_id_file(base + ".id", rc.order->ids);
            register_files(rc, base, has_id_file(rc, kind));
        }
    }
    std::lock_guard<std::mutex> lk(*rc.mu);
    *rc.write_s += since(t0);
}

void note_timers(const RunContext& rc, ggrm_t* h, double* pass_ms, double* syrk_ms) {
    double t[8] = {0};
    ggrm_get_timers(h, t);
    std::lock_guard<std::mutex> lk(*rc.mu);
    *pass_ms = std::max(*pass_ms, t[7]);
    *syrk_ms = std::max(*syrk_ms, t[3]);
}

// One GPU, a list of groups (chromosome runs) in SNP order.  Group g is written while the GPU already works on g+1.
// `part` says which piece of every group's file this handle writes (all rows, or the tiles it owns in 2D-tile mode).
void run_groups_on_device(const RunContext& rc, int device, const std::vector<int>& my_groups, bool keep_whole, ggrm_t** keep_handle,
                          double* pass_ms, double* syrk_ms, const WritePart& part, int tile_rank, int tile_world, bool report, bool write_whole_here,
                          BedRelay* relay = nullptr) {
    HandleGuard hg;
    const auto t_setup = Clock::now();
    hg.h = create_handle(rc, device, my_groups, keep_whole, tile_rank, tile_world);
    LazyPool lazy_pool(rc.cfg->writer_threads, rc.cfg->writer_threads + 2, std::max<int64_t>(kWriterBufferElems, rc.n_out));
    const auto t_push = Clock::now();
    note_phase(rc, "device setup", t_setup, t_push);
    int pending = -1;  // local index of a group that is complete on the device but not yet written
    auto group_name = [&](int li) { return ".chr" + rc.groups[static_cast<size_t>(my_groups[static_cast<size_t>(li)])].name; };
    // A chromosome that re-appears later in the BIM gives a second run with the same output name; the reference writes both, the
    // later one over the earlier (grm_work_plan.cpp:54-81, grm_engine.cpp:88-110).  The earlier file is therefore never written
    // here (its partial still feeds the whole-genome sum): same final files, and two GPUs never write one file in chromosome-shard mode.
    auto superseded = [&](int li) {
        const size_t gi = static_cast<size_t>(my_groups[static_cast<size_t>(li)]);
        for (size_t later = gi + 1; later < rc.groups.size(); ++later)
            if (rc.groups[later].name == rc.groups[gi].name) return true;
        return false;
    };
    // LOCO, this handle writes whole files: prepare the chromosome files (and the whole-genome one, written last) ahead of the writer
    std::unique_ptr<OutputPreparer> preparer;
    std::vector<int> prep_index(my_groups.size() + 1, -1);  // local group (last slot: the whole-genome file) -> preparer entry
    const bool whole_last = rc.loco && keep_whole && write_whole_here;
    if (rc.loco && rc.prefault && part.kind == WritePart::AllRows && part.create) {
        std::vector<std::vector<std::string>> bins;
        auto entry = [&](const std::string& suffix) {
            std::vector<std::string> v;
            for (int kind = 0; kind < 2; ++kind)
                if (kind == 0 ? rc.want_add : rc.want_dom) v.push_back(output_base(rc, kind, suffix) + ".bin");
            bins.push_back(std::move(v));
            return static_cast<int>(bins.size()) - 1;
        };
        for (size_t li = 0; li < my_groups.size(); ++li)
            if (!superseded(static_cast<int>(li))) prep_index[li] = entry(group_name(static_cast<int>(li)));
        if (whole_last) prep_index[my_groups.size()] = entry("");
        // with direct DMA the writer threads idle, so the page population gets their share of the cores too
        if (!bins.empty())
            preparer = std::make_unique<OutputPreparer>(std::move(bins), rc.n_out, std::max(2, rc.cfg->writer_threads) * (rc.direct_dma ? 2 : 1), 2, rc.direct_dma);
    }
    // write one prepared (or not) output entry
    auto write_entry = [&](int local_group, const std::string& suffix, size_t slot) {
        WritePart p = part;
        const int pi = prep_index[slot];
        if (preparer && pi >= 0) p.precreated = preparer->wait_ready(static_cast<size_t>(pi));
        try {
            write_outputs(rc, hg.h, local_group, suffix, p, lazy_pool.get());
        } catch (...) {
            if (preparer && pi >= 0) preparer->done(static_cast<size_t>(pi));
            throw;
        }
        if (preparer && pi >= 0) preparer->done(static_cast<size_t>(pi));
    };
    for (size_t li = 0; li < my_groups.size(); ++li) {
        const ChrRun& run = rc.groups[static_cast<size_t>(my_groups[li])];
        int64_t done = 0;
        // first chunk of this group makes the previous group fetchable
        const int64_t first_end = std::min<int64_t>(run.end, run.begin + rc.cfg->chunk_size);
        push_range(rc, hg.h, run.begin, first_end, run.end - run.begin, &done, report, relay);
        if (pending >= 0) {
            if (!superseded(pending)) write_entry(pending, group_name(pending), static_cast<size_t>(pending));
            for (int kind = 0; kind < 2; ++kind)
                if (kind == 0 ? rc.want_add : rc.want_dom) check(hg.h, ggrm_release_group(hg.h, kind, pending));
            pending = -1;
        }
        push_range(rc, hg.h, first_end, run.end, run.end - run.begin, &done, report, relay);
        if (rc.loco) pending = static_cast<int>(li);
    }
    // The matrix still accumulating when the last SNP is pushed (the whole-genome one, or the last chromosome) is read out
    // through ggrm_finish_async: its first row bands cross PCIe while the last slab is still on the tensor cores.
    const bool whole_here = !rc.loco && (write_whole_here || part.kind == WritePart::OwnedTiles);
    check(hg.h, ggrm_finish_async(hg.h));
    const auto t_readout = Clock::now();
    note_phase(rc, "push", t_push, t_readout);
    if (rc.loco && pending >= 0 && !superseded(pending)) write_entry(pending, group_name(pending), static_cast<size_t>(pending));
    if (whole_here) write_outputs(rc, hg.h, GGRM_GROUP_WHOLE, "", part, lazy_pool.get());
    check(hg.h, ggrm_finish(hg.h));
    note_timers(rc, hg.h, pass_ms, syrk_ms);
    if (rc.loco && keep_whole && (write_whole_here || part.kind == WritePart::OwnedTiles)) write_entry(GGRM_GROUP_WHOLE, "", my_groups.size());
    preparer.reset();
    const auto t_teardown = Clock::now();
    note_phase(rc, "read-out", t_readout, t_teardown);
    if (keep_handle != nullptr) {
        *keep_handle = hg.h;
        hg.h = nullptr;
    } else {
        ggrm_destroy(hg.h);
        hg.h = nullptr;
        note_phase(rc, "teardown", t_teardown, Clock::now());
    }
}

// ---------------------------------------------------------------------------------------------------------------
// SNP shards of the whole-genome matrix, band-pipelined (SURVEY 8e "fuse the reduce-scatter behind the last chunks' SYRK").
// Every GPU accumulates its SNP range; the last slabs of a pass finish row band by row band (ggrm_finish_async) and a CUDA
// host callback raises flags[rank][band] the moment a band is final on that GPU.  The 2 048-row strips of the matrix are dealt
// to the GPUs by LPT on their triangle area; the owner of a strip waits for every rank's flag of the strip's band, sums the
// peers' rows into its own over NVLink peer memory (ggrm_reduce_rows_async, read-out stream) and writes them - while all GPUs
// are still accumulating the later bands.  No collective, no barrier inside the pass.
struct Strip { int band; int64_t r0, r1; };

// Same rule as gelex_b200/sharding.py::strip_owners (compared in tests/test_sharding.py).
std::vector<std::vector<Strip>> strip_owners(int64_t n, int world, const std::vector<int64_t>& band_row_end) {
    std::vector<Strip> strips;
    for (int64_t lo = 0; lo < n; lo += kSuperRows) {
        const int64_t hi = std::min(n, lo + kSuperRows);
        int b = 0;
        while (static_cast<size_t>(b) + 1 < band_row_end.size() && band_row_end[static_cast<size_t>(b)] < hi) ++b;
        strips.push_back({b, lo, hi});
    }
    std::vector<size_t> order(strips.size());
    std::iota(order.begin(), order.end(), size_t{0});
    auto area = [&](size_t i) { return strips[i].r1 * (strips[i].r1 + 1) / 2 - strips[i].r0 * (strips[i].r0 + 1) / 2; };
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return area(a) > area(b); });
    std::vector<int64_t> load(static_cast<size_t>(world), 0);
    std::vector<int> owner(strips.size(), 0);
    for (size_t i : order) {
        const size_t r = static_cast<size_t>(std::min_element(load.begin(), load.end()) - load.begin());
        owner[i] = static_cast<int>(r);
        load[r] += area(i);
    }
    std::vector<std::vector<Strip>> out(static_cast<size_t>(world));
    for (size_t i = 0; i < strips.size(); ++i) out[static_cast<size_t>(owner[i])].push_back(strips[i]);  // row order within a rank
    return out;
}

// The same reduce-scatter at the grain that keeps every band's tail short (same rule as gelex_b200/sharding.py::piece_owners,
// compared in tests/test_sharding.py): every 2 048-row strip is cut along its columns into pieces of whole super-tile columns,
// about 64 MB of FP32 each, and the pieces of every band are dealt over ALL ranks - largest first, to the rank with the least
// work in this band, ties to the least loaded overall, then the lower rank.
struct Piece { int band; int64_t si, c0, c1; };

std::vector<std::vector<Piece>> piece_owners(int64_t n, int world, const std::vector<int64_t>& band_row_end, int64_t target_elems = int64_t{16} << 20) {
    struct P { Piece p; int64_t area; };
    std::vector<P> pieces;
    int64_t si = 0;
    for (int64_t lo = 0; lo < n; lo += kSuperRows, ++si) {
        const int64_t hi = std::min(n, lo + kSuperRows), rows = hi - lo;
        int b = 0;
        while (static_cast<size_t>(b) + 1 < band_row_end.size() && band_row_end[static_cast<size_t>(b)] < hi) ++b;
        const int64_t ncol = si + 1;
        const int64_t diag_area = hi * (hi + 1) / 2 - lo * (lo + 1) / 2 - rows * lo;
        const int64_t w = std::max<int64_t>(1, target_elems / (rows * kSuperRows));
        for (int64_t c0 = 0; c0 < ncol; c0 += w) {
            const int64_t c1 = std::min(ncol, c0 + w);
            int64_t area = 0;
            for (int64_t c = c0; c < c1; ++c) area += c < si ? rows * kSuperRows : diag_area;
            pieces.push_back({{b, si, c0, c1}, area});
        }
    }
    std::vector<int64_t> total(static_cast<size_t>(world), 0);
    std::vector<std::vector<Piece>> out(static_cast<size_t>(world));
    for (size_t b = 0; b < band_row_end.size(); ++b) {
        std::vector<size_t> idx;
        for (size_t i = 0; i < pieces.size(); ++i)
            if (pieces[i].p.band == static_cast<int>(b)) idx.push_back(i);
        std::stable_sort(idx.begin(), idx.end(), [&](size_t x, size_t y) { return pieces[x].area > pieces[y].area; });
        std::vector<int64_t> band(static_cast<size_t>(world), 0);
        for (size_t i : idx) {
            size_t best = 0;
            for (size_t q = 1; q < static_cast<size_t>(world); ++q)
                if (band[q] < band[best] || (band[q] == band[best] && total[q] < total[best])) best = q;
            band[best] += pieces[i].area;
            total[best] += pieces[i].area;
            out[best].push_back(pieces[i].p);
        }
    }
    for (auto& v : out)
        std::sort(v.begin(), v.end(), [](const Piece& a, const Piece& b) {
            return std::tie(a.band, a.si, a.c0) < std::tie(b.band, b.si, b.c0);
        });
    return out;
}

struct ShardPipe {
    std::vector<ggrm_t*> handles;                 // one per rank, created up front
    std::vector<std::pair<int64_t, int64_t>> snps;  // SNP range of each rank
    std::vector<std::vector<const void*>> accum;  // [kind][rank] device address of the accumulator
    std::vector<std::vector<Strip>> mine;         // strips each rank reduces + writes (GCTA runs: the ratio is read out by rows)
    std::vector<std::vector<Piece>> pieces;       // column pieces of strips each rank reduces + writes (every other run)
    std::vector<int32_t> flags;                   // [rank][band], written by CUDA host callbacks
    int n_bands = 1;
    std::atomic<bool> failed{false};
    int64_t snps_done = 0;                        // over all ranks (progress events; guarded by the observer's mutex)
    ProgressReplay replay;                        // ... and the second item's events of an A + D run (emit_progress_locked)
};

void run_pipelined_rank(const RunContext& rc, ShardPipe& sp, size_t r, double* pass_ms, double* syrk_ms) {
    ggrm_t* h = sp.handles[r];
    const size_t ng = sp.handles.size();
    const auto nb = static_cast<size_t>(sp.n_bands);
    check(h, ggrm_set_band_flags(h, sp.flags.data() + r * nb, 1));
    check(h, ggrm_set_push_end(h, sp.snps[r].second));  // equal slabs inside the shard
    LazyPool lazy_pool(rc.cfg->writer_threads, rc.cfg->writer_threads + 2, std::max<int64_t>(kWriterBufferElems, rc.n_out));
    BinFile files[2];
    int64_t done = 0;
    const auto t_push = Clock::now();
    push_range(rc, h, sp.snps[r].first, sp.snps[r].second, rc.m, &done, true, nullptr, &sp.snps_done, &sp.replay);
    check(h, ggrm_finish_async(h));
    const auto t0 = Clock::now();
    note_phase(rc, "push", t_push, t0);
    WriterPool& pool = lazy_pool.get();  // page-locked buffers allocated now, under the accumulation
    // opened only now: a file whose pages are being populated (Prefaulter) is opened when that has finished, which must not hold
    // up the pushes - the whole pass is in front of the first write
    for (int kind = 0; kind < 2; ++kind)
        if (kind == 0 ? rc.want_add : rc.want_dom) files[kind].open(output_base(rc, kind, "") + ".bin", rc.n_out, false, rc.writer, rc.prefault);
    auto wait_band = [&](int band) {  // every rank's rows of this band are final
        const auto t_wait = Clock::now();
        for (;;) {
            bool ready = true;
            for (size_t p = 0; p < ng && ready; ++p)
                ready = __atomic_load_n(&sp.flags[p * nb + static_cast<size_t>(band)], __ATOMIC_ACQUIRE) == 1;
            if (ready) return;
            if (sp.failed.load()) throw GelexException("multi-GPU run aborted: another GPU failed");
            if (since(t_wait) > 3600.0) throw GelexException("multi-GPU run: timed out waiting for a peer's row band");
            std::this_thread::sleep_for(std::chrono::microseconds(20));
        }
    };
    auto peers_of = [&](int kind) {
        std::vector<const void*> peers;
        for (size_t p = 0; p < ng; ++p)
            if (p != r) peers.push_back(sp.accum[static_cast<size_t>(kind)][p]);
        return peers;
    };
    try {
        // column pieces: reduce the piece's super-tiles over NVLink, then read them out one 2 048 x 2 048 super-tile at a time and
        // write their row segments at their own file offsets (as the 2D-tile writer does)
        for (const Piece& pc : rc.gcta ? std::vector<Piece>{} : sp.pieces[r]) {
            wait_band(pc.band);
            for (int kind = 0; kind < 2; ++kind) {
                if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
                const auto peers = peers_of(kind);
                for (size_t off = 0; off < peers.size(); off += 7)
                    check(h, ggrm_reduce_block_async(h, kind, GGRM_GROUP_WHOLE, peers.data() + off, static_cast<int>(std::min<size_t>(7, peers.size() - off)),
                                                     pc.si, pc.c0, pc.c1));
            }
            const int64_t row0 = pc.si * kSuperRows, rows = std::min(kSuperRows, rc.n_out - row0);
            for (int kind = 0; kind < 2; ++kind) {
                if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
                for (int64_t sc = pc.c0; sc < pc.c1; ++sc) {
                    const int64_t col0 = sc * kSuperRows, cols = std::min(kSuperRows, rc.n_out - col0);
                    float* buf = pool.acquire();
                    const int st = ggrm_fetch_block(h, kind, GGRM_GROUP_WHOLE, row0, rows, col0, cols, buf, cols);
                    if (st != GGRM_OK) {
                        pool.submit(files[kind], buf, {});
                        throw_status(st, ggrm_last_error(h));
                    }
                    pool.submit(files[kind], buf, supertile_segments(row0, col0, rows, cols));
                }
            }
        }
        bool diag_reduced = !rc.gcta_diag;
        for (const Strip& st : rc.gcta ? sp.mine[r] : std::vector<Strip>{}) {
            wait_band(st.band);
            if (!diag_reduced) {  // any band flag of a peer means its pass - and with it its diagonal sums - is complete
                reduce_gcta_diag(sp.handles, r);
                diag_reduced = true;
            }
            for (int kind = 0; kind < 2; ++kind) {
                if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
                std::vector<const void*> peers;
                for (size_t p = 0; p < ng; ++p)
                    if (p != r) peers.push_back(sp.accum[static_cast<size_t>(kind)][p]);
                for (size_t off = 0; off < peers.size(); off += 7)  // up to 7 peers per call
                    check(h, ggrm_reduce_rows_async(h, kind, GGRM_GROUP_WHOLE, peers.data() + off, static_cast<int>(std::min<size_t>(7, peers.size() - off)),
                                                    st.r0, st.r1));
            }
            for (int kind = 0; kind < 2; ++kind)  // after BOTH kinds are reduced: the GCTA ratio reads the two of them
                if (kind == 0 ? rc.want_add : rc.want_dom) write_packed_rows(h, kind, GGRM_GROUP_WHOLE, st.r0, st.r1, files[kind], pool, rc.gcta);
        }
        check(h, ggrm_finish(h));  // accumulation + read-out stream drained, device faults surfaced
    } catch (...) {
        try { pool.drain(); } catch (...) {}
        throw;
    }
    pool.drain();
    for (auto& f : files) f.close_checked();
    note_timers(rc, h, pass_ms, syrk_ms);
    note_phase(rc, "read-out", t0, Clock::now());
    std::lock_guard<std::mutex> lk(*rc.mu);
    *rc.write_s += since(t0);
}

// Rows cut on super-row (2048) boundaries into `parts` ranges of about equal lower-triangle area (the same rule as
// gelex_b200/sharding.py::row_slices, which the gloo test covers).
std::vector<std::pair<int64_t, int64_t>> row_slices(int64_t n, int parts) {
    const int64_t t = (n + 255) / 256, st = 8;
    const int64_t nsr = (t + st - 1) / st;
    std::vector<int64_t> cum(static_cast<size_t>(nsr));
    for (int64_t s = 0; s < nsr; ++s) {
        const int64_t rows = std::min(t, (s + 1) * st);
        cum[static_cast<size_t>(s)] = rows * (rows + 1) / 2;
    }
    const double total = static_cast<double>(cum.back());
    std::vector<int64_t> cuts{0};
    int64_t s = 0;
    for (int r = 1; r < parts; ++r) {
        const double target = total * r / parts;
        while (s < nsr && static_cast<double>(cum[static_cast<size_t>(s)]) < target) ++s;
        const double lo = s > 0 ? static_cast<double>(cum[static_cast<size_t>(s - 1)]) : 0.0;
        const int64_t c = (s < nsr && static_cast<double>(cum[static_cast<size_t>(s)]) - target <= target - lo) ? s + 1 : s;
        cuts.push_back(std::max(cuts.back(), std::min(nsr, c)));
    }
    cuts.push_back(nsr);
    std::vector<std::pair<int64_t, int64_t>> out;
    for (int r = 0; r < parts; ++r)
        out.emplace_back(std::min(n, cuts[static_cast<size_t>(r)] * kSuperRows), std::min(n, cuts[static_cast<size_t>(r + 1)] * kSuperRows));
    return out;
}

// Whole-genome matrix held as partial sums on several GPUs -> file.  Reduce-scatter by row slices over NVLink peer memory
// (one process owns every GPU, so peers' device pointers are addressable once peer access is on; no NCCL): GPU r sums ITS
// rows from every peer's accumulator and then writes only those rows, so the read-out uses every GPU's PCIe link and the
// file is written by several threads at once.
void reduce_scatter_and_write(const RunContext& rc, const std::vector<ggrm_t*>& handles, const std::vector<int>& devices, bool precreated) {
    const size_t ng = handles.size();
    for (size_t a = 0; a < ng; ++a)
        for (size_t b = 0; b < ng; ++b)
            if (a != b) {
                const int st = ggrm_enable_peer_access(devices[a], devices[b]);
                if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
            }
    std::vector<std::vector<const void*>> accum(2, std::vector<const void*>(ng, nullptr));
    for (int kind = 0; kind < 2; ++kind) {
        if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
        for (size_t p = 0; p < ng; ++p) {
            void* ptr = nullptr;
            check(handles[p], ggrm_accum_buffer(handles[p], kind, GGRM_GROUP_WHOLE, &ptr, nullptr, nullptr));
            accum[static_cast<size_t>(kind)][p] = ptr;
        }
    }
    if (!precreated) precreate_outputs(rc, {""});
    const auto slices = row_slices(rc.n_out, static_cast<int>(ng));
    std::vector<std::exception_ptr> errors(ng);
    std::vector<std::thread> threads;
    for (size_t r = 0; r < ng; ++r) {
        threads.emplace_back([&, r] {
            try {
                const auto [r0, r1] = slices[r];
                if (r1 <= r0) return;
                int64_t eb = 0, ee = 0;
                check(handles[r], ggrm_row_slice(handles[r], r0, r1, &eb, &ee));
                for (int kind = 0; kind < 2; ++kind) {
                    if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
                    std::vector<const void*> peers;
                    for (size_t p = 0; p < ng; ++p)
                        if (p != r) peers.push_back(accum[static_cast<size_t>(kind)][p]);
                    for (size_t off = 0; off < peers.size(); off += 7) {  // ggrm_reduce_peer takes up to 7 peers per call
                        const int cnt = static_cast<int>(std::min<size_t>(7, peers.size() - off));
                        check(handles[r], ggrm_reduce_peer(handles[r], kind, GGRM_GROUP_WHOLE, peers.data() + off, cnt, eb, ee));
                    }
                }
                if (rc.gcta_diag) reduce_gcta_diag(handles, r);
                WriterPool pool(rc.cfg->writer_threads, rc.cfg->writer_threads + 2, std::max<int64_t>(kWriterBufferElems, rc.n_out));
                WritePart part;
                part.kind = WritePart::RowSlice;
                part.r0 = r0;
                part.r1 = r1;
                part.create = false;
                write_outputs(rc, handles[r], GGRM_GROUP_WHOLE, "", part, pool);
            } catch (...) {
                errors[r] = std::current_exception();
            }
        });
    }
    for (auto& t : threads) t.join();
    for (auto& e : errors) if (e) std::rethrow_exception(e);
}

void run_multi_gpu(const RunContext& rc, bool tile_mode, double* pass_ms, double* syrk_ms) {
    const std::vector<int>& gpus = rc.cfg->gpus;
    const size_t ng = gpus.size();
    std::vector<ggrm_t*> handles(ng, nullptr);
    std::vector<std::exception_ptr> errors(ng);
    std::vector<std::thread> threads;
    struct Cleanup {  // handles go down side by side: releasing a 17 GB handle whose memory peers have mapped takes ~0.3 s
        std::vector<ggrm_t*>* hs;
        ~Cleanup() {
            std::vector<std::thread> ts;
            for (ggrm_t* h : *hs)
                if (h != nullptr) ts.emplace_back([h] { ggrm_destroy(h); });
            for (auto& t : ts) t.join();
        }
    } cleanup{&handles};
    const bool want_whole = !rc.loco || rc.cfg->write_whole;
    // function scope: these outlive the worker threads (joined below), which read them by reference
    std::vector<std::vector<int>> mine(ng);
    std::vector<int> all;
    ShardPipe pipe;
    std::unique_ptr<BedRelay> relay;  // function scope: outlives the worker threads
    std::vector<std::unique_ptr<Prefaulter>> prefaulters;  // stopped when this function returns
    bool pipelined = false, whole_precreated = false;
    if (tile_mode) {
        // 2D output tiles: every GPU streams ALL SNPs, owns disjoint 2048 x 2048 super-tiles of every matrix and writes them
        // straight to their file offsets.  No collective anywhere on the path.
        std::vector<std::string> suffixes;
        if (rc.loco) for (const ChrRun& g : rc.groups) suffixes.push_back(".chr" + g.name);
        if (want_whole) suffixes.emplace_back("");
        precreate_outputs(rc, suffixes);
        prefaulters = start_prefault(rc);
        all.resize(rc.groups.size());
        std::iota(all.begin(), all.end(), 0);
        if (rc.bed_relay) relay = std::make_unique<BedRelay>(rc, gpus, 2);
        for (size_t r = 0; r < ng; ++r) {
            threads.emplace_back([&, r] {
                try {
                    WritePart part;
                    part.kind = WritePart::OwnedTiles;
                    part.create = false;
                    run_groups_on_device(rc, gpus[r], all, rc.loco && want_whole, nullptr, pass_ms, syrk_ms, part, static_cast<int>(r), static_cast<int>(ng), r == 0, false,
                                         relay.get());
                } catch (...) {
                    if (relay) relay->fail();
                    errors[r] = std::current_exception();
                }
            });
        }
    } else if (rc.loco) {
        // chromosome runs are dealt to GPUs by LPT on their SNP counts: every partial is produced by ONE GPU, no reduce
        std::vector<size_t> order(rc.groups.size());
        std::iota(order.begin(), order.end(), size_t{0});
        std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) {
            return rc.groups[a].end - rc.groups[a].begin > rc.groups[b].end - rc.groups[b].begin;
        });
        std::vector<int64_t> load(ng, 0);
        for (size_t g : order) {
            const size_t r = static_cast<size_t>(std::min_element(load.begin(), load.end()) - load.begin());
            mine[r].push_back(static_cast<int>(g));
            load[r] += rc.groups[g].end - rc.groups[g].begin;
        }
        if (want_whole && rc.prefault) {
            // the whole-genome file is written last (reduce-scatter of the per-GPU sums): size it now, populate it during the pass
            precreate_outputs(rc, {""});
            prefaulters = start_prefault(rc, true);
            whole_precreated = true;
        }
        for (size_t r = 0; r < ng; ++r) {
            std::sort(mine[r].begin(), mine[r].end());
            threads.emplace_back([&, r] {
                try {
                    if (!mine[r].empty())
                        run_groups_on_device(rc, gpus[r], mine[r], want_whole, want_whole ? &handles[r] : nullptr, pass_ms, syrk_ms, WritePart{}, 0, 1, true, false);
                } catch (...) {
                    errors[r] = std::current_exception();
                }
            });
        }
    } else {
        // contiguous SNP ranges cut on multiples of 64; ranks whose range would be empty (tiny m) are left out
        for (size_t r = 0; r < ng; ++r) {
            const int64_t b = std::min<int64_t>(rc.m, (rc.m * static_cast<int64_t>(r) / static_cast<int64_t>(ng) + 63) / 64 * 64);
            const int64_t e = r + 1 == ng ? rc.m : std::min<int64_t>(rc.m, (rc.m * static_cast<int64_t>(r + 1) / static_cast<int64_t>(ng) + 63) / 64 * 64);
            if (e > b) pipe.snps.emplace_back(b, e);
        }
        const size_t world = pipe.snps.size();
        pipe.handles.assign(world, nullptr);
        const auto t_setup = Clock::now();
        for (size_t r = 0; r < world; ++r) {  // handles first (contexts, accumulators, rings), in parallel
            threads.emplace_back([&, r] {
                try {
                    pipe.handles[r] = create_handle(rc, gpus[r], {0}, false);
                } catch (...) {
                    errors[r] = std::current_exception();
                }
            });
        }
        for (auto& t : threads) t.join();
        threads.clear();
        for (size_t r = 0; r < world; ++r) handles[r] = pipe.handles[r];  // owned by `cleanup` from here on
        for (auto& e : errors) if (e) std::rethrow_exception(e);
        for (size_t a = 0; a < world; ++a)
            for (size_t b = 0; b < world; ++b)
                if (a != b && gpus[a] != gpus[b]) {
                    const int st = ggrm_enable_peer_access(gpus[a], gpus[b]);
                    if (st != GGRM_OK) throw_status(st, ggrm_last_error(nullptr));
                }
        pipe.accum.assign(2, std::vector<const void*>(world, nullptr));
        for (int kind = 0; kind < 2; ++kind) {
            if (!(kind == 0 ? rc.want_add : rc.want_dom)) continue;
            for (size_t r = 0; r < world; ++r) {
                void* ptr = nullptr;
                check(pipe.handles[r], ggrm_accum_address(pipe.handles[r], kind, &ptr));
                pipe.accum[static_cast<size_t>(kind)][r] = ptr;
            }
        }
        std::vector<int64_t> band_end(64, 0);
        check(pipe.handles[0], ggrm_band_layout(pipe.handles[0], band_end.data(), 64, &pipe.n_bands));
        band_end.resize(static_cast<size_t>(pipe.n_bands));
        pipe.mine = strip_owners(rc.n_out, static_cast<int>(world), band_end);
        pipe.pieces = piece_owners(rc.n_out, static_cast<int>(world), band_end);
        pipe.flags.assign(world * static_cast<size_t>(pipe.n_bands), 0);
        precreate_outputs(rc, {""});
        prefaulters = start_prefault(rc);
        note_phase(rc, "device setup", t_setup, Clock::now());
        for (size_t r = 0; r < world; ++r) {
            threads.emplace_back([&, r] {
                try {
                    run_pipelined_rank(rc, pipe, r, pass_ms, syrk_ms);
                } catch (...) {
                    pipe.failed.store(true);
                    errors[r] = std::current_exception();
                }
            });
        }
        pipelined = true;
    }
    for (auto& t : threads) t.join();
    if (relay) {
        // an uploader's own failure (unreadable BED, device error) is the cause; the ranks it starved only report "aborted"
        try {
            relay->stop_and_rethrow();
        } catch (...) {
            relay.reset();
            throw;
        }
        relay.reset();
    }
    for (auto& e : errors) if (e) std::rethrow_exception(e);
    if (want_whole && !tile_mode && !pipelined) {
        std::vector<ggrm_t*> live;
        std::vector<int> devs;
        for (size_t r = 0; r < ng; ++r)
            if (handles[r] != nullptr) {
                live.push_back(handles[r]);
                devs.push_back(gpus[r]);
            }
        if (!live.empty()) reduce_scatter_and_write(rc, live, devs, whole_precreated);
    }
}

}  // namespace

int GrmEngine::gpus_worth_using(int64_t n, int64_t m, int kinds, int precision, int offered, int n_outputs) {
    if (offered <= 1 || n <= 0 || m <= 0) return std::max(1, offered);
    // measured on the 2- and 8-GPU boxes (profiles/r02_summary.md sections 6 and 9)
    constexpr double kContextSeconds = 0.30;    // per GPU, serialised by the driver: 2.3 s for eight
    constexpr double kMultiGpuSeconds = 0.40;   // reduce-scatter + piece-grain read-out of a run on more than one GPU
    const double rate = precision == GGRM_PREC_FP64 ? 3.0e13 : 1.3e15;
    const double t1 = static_cast<double>(n) * (static_cast<double>(n) + 1.0) * static_cast<double>(m) * std::max(1, kinds) / rate;
    int best = 1;
    double best_cost = kContextSeconds + t1;
    for (int k = 2; k <= offered; ++k) {
        const double cost = kContextSeconds * k + t1 / k + kMultiGpuSeconds;
        if (cost < best_cost) {
            best_cost = cost;
            best = k;
        }
    }
    // write-bound runs (a LOCO run's 23 x 5 GB): a second GPU brings a second PCIe link and writer pool (C4: 17.7 -> 14.9 s), more do not
    const double out_bytes = 2.0 * static_cast<double>(n) * static_cast<double>(n) * std::max(1, n_outputs);
    if (best == 1 && out_bytes >= 32e9) best = 2;
    // ... and never fewer than the FP32 (FP64: double) triangle needs: n^2 / 2 entries per matrix, ~120 GB of a B200 for accumulators
    const double acc_bytes = 0.5 * static_cast<double>(n) * static_cast<double>(n) * (precision == GGRM_PREC_FP64 ? 8.0 : 4.0) * std::max(1, kinds);
    const int floor_k = static_cast<int>(std::ceil(acc_bytes / 120e9));
    return std::min(offered, std::max(best, floor_k));
}

std::vector<std::pair<int64_t, int64_t>> GrmEngine::row_slices_for(int64_t n, int parts) { return row_slices(n, parts); }

std::vector<GrmEngine::OwnedPiece> GrmEngine::piece_owners_for(int64_t n, int world, const std::vector<int64_t>& band_row_end) {
    std::vector<OwnedPiece> out;
    const auto per_rank = piece_owners(n, world, band_row_end);
    for (size_t r = 0; r < per_rank.size(); ++r)
        for (const Piece& pc : per_rank[r]) out.push_back({static_cast<int>(r), pc.band, pc.si, pc.c0, pc.c1});
    return out;
}

std::vector<GrmEngine::OwnedStrip> GrmEngine::strip_owners_for(int64_t n, int world, const std::vector<int64_t>& band_row_end) {
    std::vector<OwnedStrip> out;
    const auto per_rank = strip_owners(n, world, band_row_end);
    for (size_t r = 0; r < per_rank.size(); ++r)
        for (const Strip& st : per_rank[r]) out.push_back({static_cast<int>(r), st.band, st.r0, st.r1});
    return out;
}

namespace {

WriterMode resolve_writer_mode(WriterMode configured) {
    if (configured != WriterMode::Auto) return configured;
    const char* e = std::getenv("GELEX_B200_WRITER");
    if (e == nullptr) return WriterMode::Auto;
    const std::string v = lower(e);
    if (v == "pwrite") return WriterMode::Pwrite;
    if (v == "mmap") return WriterMode::Mmap;
    if (v == "auto" || v.empty()) return WriterMode::Auto;
    throw ArgumentValidationException("GELEX_B200_WRITER must be auto, pwrite or mmap");
}

// entries (i, j0 .. j0 + len) of the self-test matrix: float((31 i + j) mod 65521) + 0.5, copied from two periods of the pattern
// (the producer must not be the bottleneck of a writer measurement)
void selftest_fill(float* out, int64_t i, int64_t j0, int64_t len) {
    constexpr int64_t kP = 65521;
    static const std::vector<float> table = [] {
        std::vector<float> t(static_cast<size_t>(2 * kP));
        for (int64_t k = 0; k < 2 * kP; ++k) t[static_cast<size_t>(k)] = static_cast<float>(k % kP) + 0.5f;
        return t;
    }();
    while (len > 0) {
        const int64_t start = (i * 31 + j0) % kP;
        const int64_t take = std::min(len, kP);
        std::memcpy(out, table.data() + start, static_cast<size_t>(take) * sizeof(float));
        out += take;
        j0 += take;
        len -= take;
    }
}

}  // namespace

double GrmEngine::writer_selftest(const fs::path& bin_path, int64_t n, bool tiles, int parts, int writer_threads, WriterMode mode, bool* mapped,
                                  int prefault_threads, double* prefault_seconds) {
    if (n < 1 || parts < 1 || writer_threads < 1) throw ArgumentValidationException("writer_selftest: invalid argument");
    mode = resolve_writer_mode(mode);
    if (mode != WriterMode::Pwrite && !file_system_has_room(bin_path, tri(n) * 4)) mode = WriterMode::Pwrite;
    {
        BinFile f;
        f.open(bin_path, n, true);
        f.close_checked();
    }
    // what the engine does while the GPUs accumulate (Config::prefault): here it runs to completion first and is timed on its own
    if (prefault_seconds != nullptr) *prefault_seconds = 0.0;
    std::unique_ptr<Prefaulter> prefaulter;
    if (prefault_threads > 0 && mode != WriterMode::Pwrite) {
        const auto tp = Clock::now();
        prefaulter = std::make_unique<Prefaulter>(bin_path.string(), prefault_threads);
        prefaulter->wait();
        if (prefault_seconds != nullptr) *prefault_seconds = since(tp);
    }
    // the work of each part, as the engine deals it: row slices (SNP-shard writers) or super-tiles round-robin (2D tiles)
    const auto slices = row_slices(n, parts);
    std::vector<std::pair<int64_t, int64_t>> supers;
    for (int64_t si = 0; si * kSuperRows < n; ++si)
        for (int64_t sj = 0; sj <= si; ++sj) supers.emplace_back(si, sj);
    std::vector<std::exception_ptr> errors(static_cast<size_t>(parts));
    std::vector<int> any_map(static_cast<size_t>(parts), 0);
    std::vector<std::thread> threads;
    const auto t0 = Clock::now();
    for (int r = 0; r < parts; ++r) {
        threads.emplace_back([&, r] {
            try {
                WriterPool pool(writer_threads, writer_threads + 2, std::max<int64_t>(kWriterBufferElems, n));
                BinFile f;
                f.open(bin_path, n, false, mode);
                any_map[static_cast<size_t>(r)] = f.map != nullptr ? 1 : 0;
                if (!tiles) {
                    int64_t row = slices[static_cast<size_t>(r)].first;
                    const int64_t r1 = slices[static_cast<size_t>(r)].second;
                    while (row < r1) {
                        const int64_t cnt = rows_fitting(row, r1, pool.buffer_elems());
                        float* buf = pool.acquire();
                        float* o = buf;
                        for (int64_t i = row; i < row + cnt; o += i + 1, ++i) selftest_fill(o, i, 0, i + 1);
                        pool.submit(f, buf, {Segment{0, tri(row) * 4, (tri(row + cnt) - tri(row)) * 4}});
                        row += cnt;
                    }
                } else {
                    for (size_t t = static_cast<size_t>(r); t < supers.size(); t += static_cast<size_t>(parts)) {
                        const int64_t row0 = supers[t].first * kSuperRows, col0 = supers[t].second * kSuperRows;
                        const int64_t rows = std::min(kSuperRows, n - row0), cols = std::min(kSuperRows, n - col0);
                        float* buf = pool.acquire();
                        for (int64_t i = 0; i < rows; ++i) selftest_fill(buf + i * cols, row0 + i, col0, cols);
                        pool.submit(f, buf, supertile_segments(row0, col0, rows, cols));
                    }
                }
                pool.drain();
                f.close_checked();
            } catch (...) {
                errors[static_cast<size_t>(r)] = std::current_exception();
            }
        });
    }
    for (auto& t : threads) t.join();
    const double dt = since(t0);
    for (auto& e : errors) if (e) std::rethrow_exception(e);
    if (mapped != nullptr) *mapped = std::any_of(any_map.begin(), any_map.end(), [](int v) { return v != 0; });
    return dt;
}

int64_t GrmEngine::relay_selftest(int n_chunks, int world, int slots, int chunk_bytes, int fail_at_chunk) {
    if (n_chunks < 0 || world < 1 || slots < 1 || chunk_bytes < 1) throw ArgumentValidationException("relay_selftest: invalid argument");
    const auto pattern = [](size_t j, size_t k) { return static_cast<uint8_t>((j * 131 + k * 7 + (j >> 3)) & 0xff); };
    std::vector<std::vector<uint8_t>> store(static_cast<size_t>(world) * static_cast<size_t>(slots), std::vector<uint8_t>(static_cast<size_t>(chunk_bytes)));
    auto slot_mem = [&](size_t owner, int slot) -> std::vector<uint8_t>& { return store[owner * static_cast<size_t>(slots) + static_cast<size_t>(slot)]; };
    ChunkRelay relay(static_cast<size_t>(n_chunks), static_cast<size_t>(world), static_cast<size_t>(slots), [&](size_t owner, int slot, size_t j) {
        if (fail_at_chunk >= 0 && j == static_cast<size_t>(fail_at_chunk)) throw FileFormatException("relay_selftest: injected upload failure");
        auto& m = slot_mem(owner, slot);
        if ((j * 2654435761u >> 7) % 5 == 0) std::this_thread::sleep_for(std::chrono::microseconds(200));  // uneven uploaders
        for (size_t k = 0; k < m.size(); ++k) m[k] = pattern(j, k);
    });
    std::atomic<int64_t> bad{0};
    std::vector<std::exception_ptr> errors(static_cast<size_t>(world));
    std::vector<std::thread> consumers;
    for (int r = 0; r < world; ++r) {
        consumers.emplace_back([&, r] {
            try {
                for (size_t j = 0; j < static_cast<size_t>(n_chunks); ++j) {
                    relay.acquire(j);
                    const auto& m = slot_mem(relay.owner_of(j), relay.slot_of(j));
                    if ((j + static_cast<size_t>(r)) % 7 == 0) std::this_thread::sleep_for(std::chrono::microseconds(150));  // uneven consumers
                    int64_t wrong = 0;
                    for (size_t k = 0; k < m.size(); ++k) wrong += m[k] != pattern(j, k);
                    if (wrong != 0) bad.fetch_add(1);
                    relay.release(j);
                }
            } catch (...) {
                relay.fail();
                errors[static_cast<size_t>(r)] = std::current_exception();
            }
        });
    }
    for (auto& t : consumers) t.join();
    relay.stop_and_rethrow();  // an uploader's failure is the cause ...
    for (auto& e : errors) if (e) std::rethrow_exception(e);  // ... otherwise a consumer's
    return bad.load();
}

void GrmEngine::compute(const GrmObserver& observer) {
    const auto t_start = Clock::now();
    summary_ = Summary{};
    if (config_.gpus.empty()) throw ArgumentValidationException("at least one GPU is required (there is no CPU fallback)");
    if (config_.gcta && (config_.mode != GrmType::A || config_.do_loco))
        throw ArgumentValidationException("--gcta writes the additive matrix of the whole genome: use --add without --dom / --loco");
    if (config_.gcta_diag && (!config_.gcta || config_.method != GenotypeProcessMethod::StandardizeHWE))
        throw ArgumentValidationException("--gcta-diag is GCTA's diagonal for its own standardisation: use it with --gcta and --method StandardizeHWE");
    fs::path fam = config_.bed_path, bim = config_.bed_path;
    fam.replace_extension(".fam");
    bim.replace_extension(".bim");

    // CUDA contexts and kernel images come up on helper threads while this thread parses the metadata (0.3-0.5 s per cold
    // process otherwise in front of the first push).  Failures are not reported here: ggrm_create repeats the checks and owns the
    // error, after every input has been validated in the reference's order.
    struct Warmers {
        std::vector<std::thread> t;
        ~Warmers() { for (auto& x : t) if (x.joinable()) x.join(); }
    } warmers;
    for (int dev : config_.gpus) warmers.t.emplace_back([dev] { (void)ggrm_warm_device(dev); });

    // GRM::GRM(bed_path): SampleManager::create_finalized + BedPipe (grm.cpp:36-40, bed_pipe.cpp:69-98)
    const std::vector<std::string> raw_ids = load_fam_ids(fam);
    const SampleOrder order = make_sample_order(raw_ids);
    const int64_t m = static_cast<int64_t>(count_total_lines(bim));
    const int64_t n_raw = static_cast<int64_t>(raw_ids.size());
    const int64_t bpv = (n_raw + 3) / 4;
    const BedMapping bed(config_.bed_path, m, bpv);
    if (observer) observer(GrmDataLoadedEvent{order.ids.size(), static_cast<size_t>(m)});
    const auto t_meta1 = Clock::now();

    const std::vector<std::string> chroms = load_bim_chromosomes(bim);
    std::string pattern_suffix;
    size_t n_groups = 0;
    const std::vector<GrmWorkItem> items = build_plan(chroms, config_.mode, config_.do_loco, &pattern_suffix, &n_groups);
    size_t total_work = 0;
    for (const auto& it : items)
        for (const auto& r : it.ranges) total_work += static_cast<size_t>(r.second - r.first);
    if (observer) observer(GrmComputeStartedEvent{total_work});

    std::mutex mu;
    std::vector<std::string> files;
    double write_s = 0, pass_ms = 0, syrk_ms = 0;
    RunContext rc{};
    rc.cfg = &config_;
    rc.order = &order;
    rc.bed = &bed;
    rc.n_raw = n_raw;
    rc.n_out = static_cast<int64_t>(order.ids.size());
    rc.m = m;
    rc.bpv = bpv;
    rc.loco = config_.do_loco;
    rc.groups = config_.do_loco ? chromosome_runs(chroms) : std::vector<ChrRun>{{"", 0, static_cast<int64_t>(chroms.size())}};
    rc.gcta = config_.gcta;
    rc.gcta_diag = config_.gcta && config_.gcta_diag;
    rc.want_add = config_.mode != GrmType::D;
    rc.want_dom = config_.mode != GrmType::A || config_.gcta;
    // FP16 operands overflow when a standardised value exceeds 65 504.  The scaled orthogonal dominance codes reach |z| ~ n (a
    // homozygote of an allele seen once: -q/p = -(n - 1)), so from 65 505 samples on such a run takes BF16 operands (FP32 exponent
    // range) instead; anything else that overflows is caught on the device (ggrm_finish fails with "FP16 operand overflow").
    rc.precision = config_.precision;
    if (config_.precision == GGRM_PREC_FP16 && rc.want_dom && !config_.gcta && rc.n_out > 65504 &&
        (config_.method == GenotypeProcessMethod::OrthStandardizeHWE || config_.method == GenotypeProcessMethod::OrthStandardize))
        rc.precision = GGRM_PREC_BF16;
    summary_.precision_used = rc.precision;
    rc.observer = &observer;
    rc.mu = &mu;
    rc.files = &files;
    rc.write_s = &write_s;
    rc.t_start = t_start;
    rc.phases = &summary_.phases;
    note_phase(rc, "metadata", t_start, t_meta1);
    note_phase(rc, "work plan", t_meta1, Clock::now());

    const int n_gpus = static_cast<int>(config_.gpus.size());
    if (config_.writer_threads < 1) throw ArgumentValidationException("writer_threads must be >= 1");
    // The CLI refuses chunk_size <= 0 (grm_config.cpp:42-46).  The reference ENGINE does not look at it until the first
    // BedPipe::load_chunk(start, start + chunk_size), whose range check fails (bed_pipe.cpp:38-52) - after every input has been
    // validated.  Same here, before any loop steps by it.
    if (config_.chunk_size <= 0 && !rc.groups.empty() && !chroms.empty()) {
        const int64_t s0 = rc.groups.front().begin;
        throw ColumnRangeException("invalid chunk range: [" + std::to_string(s0) + ", " + std::to_string(s0 + config_.chunk_size) +
                                   "). Total SNPs: " + std::to_string(m));
    }
    rc.writer = resolve_writer_mode(config_.writer);
    {
        const char* e = std::getenv("GELEX_B200_BED_RELAY");
        rc.bed_relay = config_.bed_relay || (e != nullptr && e[0] == '1');
        // populate output pages ahead of the writers: on by default (measured on the GPU box: C2 read-out + write 1.12 -> 0.42 s,
        // profiles/r02a_cli_writers.jsonl); GELEX_B200_PREFAULT=0 turns it off
        const char* pf = std::getenv("GELEX_B200_PREFAULT");
        rc.prefault = (pf != nullptr ? pf[0] != '0' : config_.prefault) && rc.writer != WriterMode::Pwrite;  // a forced pwrite stays pwrite
        // Opt-in (GELEX_B200_DIRECT_DMA=1): measured on the GPU box, page-locking a 5 GB mapping (0.4 s) holds the driver lock
        // the pushing thread needs - C2 read-out 0.42 -> 0.35 s but accumulate 1.05 -> 1.42 s, C4 17.7 -> 34.9 s
        // (profiles/r02_summary.md) - so the staged writer stays the default
        const char* dd = std::getenv("GELEX_B200_DIRECT_DMA");
        rc.direct_dma = rc.prefault && dd != nullptr && dd[0] == '1';
    }
    if (rc.writer != WriterMode::Pwrite) {
        // every .bin of the run: one per work item, plus the whole-genome sums of a LOCO run, or the GCTA pair
        int64_t bins = static_cast<int64_t>(items.size());
        if (config_.do_loco && config_.write_whole) bins += (rc.want_add ? 1 : 0) + (rc.want_dom ? 1 : 0);
        if (config_.gcta) bins = 2;
        if (!file_system_has_room(config_.out_prefix, bins * tri(rc.n_out) * 4)) rc.writer = WriterMode::Pwrite;
    }
    summary_.outputs_mapped = rc.writer != WriterMode::Pwrite;
    if (items.empty() || rc.groups.empty() || chroms.empty()) {
        // nothing to accumulate (empty BIM is rejected earlier as "is empty")
    } else if (n_gpus == 1) {
        std::vector<int> all(rc.groups.size());
        std::iota(all.begin(), all.end(), 0);
        // one GPU writes everything itself, the whole-genome matrix included (non-LOCO: the only output; LOCO: with write_whole)
        WritePart part;
        std::vector<std::unique_ptr<Prefaulter>> prefaulters;
        if (rc.prefault && !rc.loco) {
            precreate_outputs(rc, {""});  // the file exists from the start so that its pages can be populated during the pass
            prefaulters = start_prefault(rc);
            part.create = false;
        }
        run_groups_on_device(rc, config_.gpus[0], all, config_.do_loco && config_.write_whole, nullptr, &pass_ms, &syrk_ms, part, 0, 1, true,
                             !config_.do_loco || config_.write_whole);
    } else {
        // SNP shards (+ one reduce) while the FP32 triangle fits comfortably per GPU, 2D output tiles from n >= 100 000
        const bool tile_mode = config_.partition == Partition::Tile || (config_.partition == Partition::Auto && rc.n_out >= 100000);
        run_multi_gpu(rc, tile_mode, &pass_ms, &syrk_ms);
    }

    if (observer) {
        observer(GrmProgressEvent{total_work, total_work, true});
        observer(GrmFilesWrittenEvent{config_.gcta ? files.size() : items.size() * 2, fs::absolute(fs::path(config_.out_prefix)).parent_path().string(),
                                      config_.out_prefix + (config_.gcta ? std::string(".grm.{bin|N.bin|id}") : pattern_suffix)});
    }
    summary_.num_samples = order.ids.size();
    summary_.num_snps = static_cast<size_t>(m);
    summary_.files = files;
    summary_.num_files = files.size();
    summary_.write_s = write_s;
    summary_.device_pass_ms = pass_ms;
    summary_.syrk_ms = syrk_ms;
    summary_.wall_s = since(t_start);
}

}  // namespace gelex::b200

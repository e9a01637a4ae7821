// Read-only LMDB (liblmdb 0.9 on-disk format, data version 1) reader + batched record gather.
//
// Stands in for py-lmdb under the reference's LMDBReader (torchbooster/lmdb.py:48-83:
// lmdb.open(readonly=True, lock=False, readahead=False) -> txn.get(key)), which is absent from
// this image.  Only what that call path needs: newest valid meta page, main DB B-tree descent
// with memcmp key order, leaf nodes and overflow ("big data") values.  No writes, no locks, no
// sub-databases / dupsort / reverse or integer keys (refused at open).  Values are returned as pointers into
// the private read-only map.
//
// On-disk layout restated from liblmdb's mdb.c (not vendored in the reference; no .mdb fixture
// exists there either -- parity for this reader is UNPINNED):
//   page header (16 B): pgno u64 | pad u16 | flags u16 | lower u16 | upper u16   (overflow: pages u32)
//   flags: BRANCH 0x01, LEAF 0x02, OVERFLOW 0x04, META 0x08
//   node ptrs: u16 offsets from the page start, beginning at byte 16; count = (lower - 16) / 2
//   node (8 B + key + data): lo u16 | hi u16 | flags u16 | ksize u16 | key | data
//       leaf  : data size = lo | hi << 16 ; flags & 0x01 (BIGDATA) -> data is the u64 pgno of an overflow run
//       branch: child pgno = lo | hi << 16 | flags << 32 ; node 0 carries no key
//   meta (at byte 16 of pages 0 and 1): magic 0xBEEFC0DE u32 | version u32 | address u64 | mapsize u64 |
//       dbs[2] x 48 B {pad u32, flags u16, depth u16, branch u64, leaf u64, overflow u64, entries u64, root u64} |
//       last_pg u64 | txnid u64 ; dbs[0].pad = page size ; dbs[1] = main DB
#include <errno.h>
#include <fcntl.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "tb_common.cuh"

struct tb_lmdb {
    const uint8_t* map;
    size_t size;
    uint32_t psize;
    uint64_t root, entries, last_pg;
    uint16_t depth, flags;
};

namespace {

const uint32_t kMagic = 0xBEEFC0DE;
const uint64_t kInvalid = ~(uint64_t)0;
enum { P_BRANCH = 0x01, P_LEAF = 0x02, P_OVERFLOW = 0x04, P_META = 0x08, P_LEAF2 = 0x20 };
enum { F_BIGDATA = 0x01, F_SUBDATA = 0x02, F_DUPDATA = 0x04 };

template <typename T>
inline T rd(const uint8_t* p) {
    T v;
    memcpy(&v, p, sizeof(T));
    return v;
}

struct Meta {
    bool ok;
    uint32_t psize;
    uint64_t root, entries, last_pg, txnid;
    uint16_t depth, flags;
};

Meta read_meta(const uint8_t* page, size_t avail) {
    Meta m{};
    if (avail < 16 + 136) return m;
    if (!(rd<uint16_t>(page + 10) & P_META)) return m;
    const uint8_t* mm = page + 16;
    if (rd<uint32_t>(mm) != kMagic || rd<uint32_t>(mm + 4) != 1) return m;
    m.psize = rd<uint32_t>(mm + 24);               // dbs[0].pad
    const uint8_t* main = mm + 24 + 48;            // dbs[1]
    m.flags = rd<uint16_t>(main + 4);
    m.depth = rd<uint16_t>(main + 6);
    m.entries = rd<uint64_t>(main + 32);
    m.root = rd<uint64_t>(main + 40);
    m.last_pg = rd<uint64_t>(mm + 24 + 96);
    m.txnid = rd<uint64_t>(mm + 24 + 104);
    m.ok = m.psize >= 512 && (m.psize & (m.psize - 1)) == 0;
    return m;
}

inline int cmp_key(const void* a, size_t an, const void* b, size_t bn) {
    const int c = memcmp(a, b, an < bn ? an : bn);
    if (c) return c;
    return an < bn ? -1 : (an > bn ? 1 : 0);
}

// 0 found, 1 absent, <0 corrupt
int lookup(const tb_lmdb* db, const void* key, size_t klen, const uint8_t** data, size_t* dlen) {
    if (db->root == kInvalid) return 1;
    uint64_t pgno = db->root;
    for (int level = 0; level < 64; ++level) {
        if (pgno > db->last_pg || (pgno + 1) * (uint64_t)db->psize > db->size) return -1;
        const uint8_t* pg = db->map + pgno * db->psize;
        const uint16_t fl = rd<uint16_t>(pg + 10);
        const uint16_t lower = rd<uint16_t>(pg + 12);
        if (lower < 16 || lower > db->psize) return -1;
        const int n = (lower - 16) >> 1;
        auto node = [&](int i) -> const uint8_t* {
            const uint16_t off = rd<uint16_t>(pg + 16 + 2 * i);
            return (off >= 16 && (uint32_t)off + 8 <= db->psize) ? pg + off : nullptr;
        };
        if (fl & P_BRANCH) {
            if (n < 1) return -1;
            // last i >= 1 whose key <= search key, else 0 (node 0 has an implicit lowest key)
            int lo = 1, hi = n - 1, pick = 0;
            while (lo <= hi) {
                const int mid = (lo + hi) >> 1;
                const uint8_t* nd = node(mid);
                if (!nd) return -1;
                const uint16_t ks = rd<uint16_t>(nd + 6);
                if ((size_t)(nd - pg) + 8 + ks > db->psize) return -1;      // key runs off the page: corrupt
                if (cmp_key(nd + 8, ks, key, klen) <= 0) { pick = mid; lo = mid + 1; }
                else hi = mid - 1;
            }
            const uint8_t* nd = node(pick);
            if (!nd) return -1;
            pgno = (uint64_t)rd<uint16_t>(nd) | ((uint64_t)rd<uint16_t>(nd + 2) << 16) | ((uint64_t)rd<uint16_t>(nd + 4) << 32);
            continue;
        }
        if (!(fl & P_LEAF) || (fl & P_LEAF2)) return -1;
        int lo = 0, hi = n - 1;
        while (lo <= hi) {
            const int mid = (lo + hi) >> 1;
            const uint8_t* nd = node(mid);
            if (!nd) return -1;
            const uint16_t ks = rd<uint16_t>(nd + 6);
            if ((size_t)(nd - pg) + 8 + ks > db->psize) return -1;
            const int c = cmp_key(nd + 8, ks, key, klen);
            if (c == 0) {
                const uint16_t nf = rd<uint16_t>(nd + 4);
                if (nf & (F_SUBDATA | F_DUPDATA)) return -1;
                const size_t ds = (size_t)rd<uint16_t>(nd) | ((size_t)rd<uint16_t>(nd + 2) << 16);
                const uint8_t* d = nd + 8 + ks;
                if (nf & F_BIGDATA) {
                    const uint64_t opg = rd<uint64_t>(d);
                    if (opg > db->last_pg) return -1;
                    const uint8_t* op = db->map + opg * db->psize;
                    if ((size_t)(op - db->map) + 16 + ds > db->size) return -1;
                    if (!(rd<uint16_t>(op + 10) & P_OVERFLOW)) return -1;
                    d = op + 16;
                } else if ((size_t)(d - pg) + ds > db->psize) {
                    return -1;
                }
                *data = d;
                *dlen = ds;
                return 0;
            }
            if (c < 0) lo = mid + 1;
            else hi = mid - 1;
        }
        return 1;
    }
    return -1;
}

}  // namespace

// path: an LMDB directory (data.mdb inside, py-lmdb's default subdir=True) or the data file.
extern "C" int tb_lmdb_open(const char* path, tb_lmdb** out) {
    TB_REQUIRE(path && out, "tb_lmdb_open: NULL argument");
    std::string file(path);
    struct stat st;
    if (stat(file.c_str(), &st) != 0) return tb_fail(TB_ENOENT, "tb_lmdb_open: %s: %s", path, strerror(errno));
    if (S_ISDIR(st.st_mode)) {
        file += "/data.mdb";
        if (stat(file.c_str(), &st) != 0) return tb_fail(TB_ENOENT, "tb_lmdb_open: %s: %s", file.c_str(), strerror(errno));
    }
    const int fd = open(file.c_str(), O_RDONLY | O_CLOEXEC);
    if (fd < 0) return tb_fail(TB_ENOENT, "tb_lmdb_open: %s: %s", file.c_str(), strerror(errno));
    const size_t size = (size_t)st.st_size;
    if (size < 2 * 512) {
        close(fd);
        return tb_fail(TB_EINVAL, "tb_lmdb_open: %s is too small to be an LMDB data file", file.c_str());
    }
    // Read-only private mapping, no swap reservation: what lmdb.open(readonly=True, lock=False) gives the
    // reference (torchbooster/lmdb.py:50-58) -- a database larger than RAM + swap opens, and nothing is charged
    // against the commit limit.  Only tb_lmdb_cuda_register needs more (it flips the protection there).
    void* map = mmap(nullptr, size, PROT_READ, MAP_PRIVATE | MAP_NORESERVE, fd, 0);
    close(fd);
    if (map == MAP_FAILED) return tb_fail(TB_EINVAL, "tb_lmdb_open: mmap(%s): %s", file.c_str(), strerror(errno));
    // readahead=False in the reference: records are read at random
    madvise(map, size, MADV_RANDOM);
    const uint8_t* base = (const uint8_t*)map;
    Meta m0 = read_meta(base, size);
    Meta m1{};
    if (m0.ok && (size_t)m0.psize * 2 <= size) m1 = read_meta(base + m0.psize, size - m0.psize);
    if (!m0.ok && !m1.ok) {
        // page size unknown without meta 0: probe the usual sizes for meta 1
        for (uint32_t ps = 512; ps <= 65536 && !m1.ok; ps <<= 1)
            if ((size_t)ps * 2 <= size) {
                m1 = read_meta(base + ps, size - ps);
                if (m1.ok && m1.psize != ps) m1.ok = false;
            }
    }
    if (!m0.ok && !m1.ok) {
        munmap(map, size);
        return tb_fail(TB_EINVAL, "tb_lmdb_open: %s has no valid LMDB meta page (magic 0xBEEFC0DE, version 1)", file.c_str());
    }
    const Meta& m = (!m1.ok || (m0.ok && m0.txnid >= m1.txnid)) ? m0 : m1;
    // Main-DB flags change the key order or the node format (MDB_REVERSEKEY 0x02, MDB_DUPSORT 0x04,
    // MDB_INTEGERKEY 0x08, MDB_DUPFIXED 0x10, MDB_INTEGERDUP 0x20, MDB_REVERSEDUP 0x40): the descent below
    // assumes plain memcmp order and single values, so such a database is refused instead of misread.
    if (m.flags & 0x7e) {
        munmap(map, size);
        return tb_fail(TB_ENOTSUP, "tb_lmdb_open: %s: main database flags 0x%x (reverse / integer keys or duplicates) are not supported",
                       file.c_str(), (unsigned)m.flags);
    }
    tb_lmdb* db = new tb_lmdb;
    db->map = base; db->size = size; db->psize = m.psize; db->root = m.root; db->entries = m.entries;
    db->last_pg = m.last_pg; db->depth = m.depth; db->flags = m.flags;
    *out = db;
    return TB_OK;
}

extern "C" int tb_lmdb_cuda_unregister(tb_lmdb* db);

extern "C" int tb_lmdb_close(tb_lmdb* db) {
    if (!db) return TB_OK;
    tb_lmdb_cuda_unregister(db);
    munmap((void*)db->map, db->size);
    delete db;
    return TB_OK;
}

extern "C" int64_t tb_lmdb_entries(const tb_lmdb* db) { return db ? (int64_t)db->entries : 0; }

// 0 and (*data, *len) on success; TB_ENOENT when the key is absent.
extern "C" int tb_lmdb_get(const tb_lmdb* db, const void* key, int64_t klen, const void** data, int64_t* len) {
    TB_REQUIRE(db && key && data && len && klen >= 0, "tb_lmdb_get: bad argument");
    const uint8_t* d = nullptr;
    size_t n = 0;
    const int rc = lookup(db, key, (size_t)klen, &d, &n);
    if (rc == 1) return tb_fail(TB_ENOENT, "tb_lmdb_get: key not found");
    if (rc < 0) return tb_fail(TB_EINVAL, "tb_lmdb_get: corrupt or unsupported page structure");
    *data = d;
    *len = (int64_t)n;
    return TB_OK;
}

// Batched form of `reader[idx]` (torchbooster/lmdb.py:96-97: key = str(idx)): copies n fixed-size
// records into one contiguous (ideally pinned) host buffer, dst + i*record_bytes, on `threads`
// host threads.  Every record must be exactly record_bytes long.
extern "C" int tb_lmdb_gather(const tb_lmdb* db, const int64_t* indices, int64_t n, int64_t record_bytes,
                                     void* dst, int threads) {
    TB_REQUIRE(db && (n == 0 || (indices && dst)) && n >= 0 && record_bytes >= 0, "tb_lmdb_gather: bad argument");
    if (threads < 1) threads = 1;
    if (threads > n) threads = (int)(n ? n : 1);
    std::vector<int> status(threads, 0);
    std::vector<int64_t> bad(threads, -1);
    auto work = [&](int t) {
        char key[24];
        for (int64_t i = t; i < n; i += threads) {
            const int kl = snprintf(key, sizeof(key), "%lld", (long long)indices[i]);
            const uint8_t* d = nullptr;
            size_t len = 0;
            const int rc = lookup(db, key, (size_t)kl, &d, &len);
            if (rc != 0 || (int64_t)len != record_bytes) {
                status[t] = rc == 1 ? 1 : (rc < 0 ? 2 : 3);
                bad[t] = indices[i];
                return;
            }
            memcpy((uint8_t*)dst + i * record_bytes, d, len);
        }
    };
    if (threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto& th : pool) th.join();
    }
    for (int t = 0; t < threads; ++t) {
        if (status[t] == 1) return tb_fail(TB_ENOENT, "tb_lmdb_gather: key %lld not found", (long long)bad[t]);
        if (status[t] == 2) return tb_fail(TB_EINVAL, "tb_lmdb_gather: corrupt page structure at key %lld", (long long)bad[t]);
        if (status[t] == 3) return tb_fail(TB_EINVAL, "tb_lmdb_gather: record %lld is not %lld bytes", (long long)bad[t], (long long)record_bytes);
    }
    return TB_OK;
}

// Records of different sizes (images stored with a small header, torchbooster_b200/lmdb.py
// write_ragged_image_lmdb).  Two uses of one call: with dst == NULL it reports sizes[i] = length of
// record str(indices[i]); with dst it also copies record i to dst + offsets[i] (the caller laid the
// offsets out from the sizes of a first call) and checks every record still has that size.
extern "C" int tb_lmdb_gather_ragged(const tb_lmdb* db, const int64_t* indices, int64_t n, int64_t* sizes,
                                            const int64_t* offsets, void* dst, int64_t capacity, int threads) {
    TB_REQUIRE(db && n >= 0 && (n == 0 || (indices && sizes)), "tb_lmdb_gather_ragged: bad argument");
    TB_REQUIRE(dst == nullptr || offsets != nullptr, "tb_lmdb_gather_ragged: a destination needs offsets");
    if (threads < 1) threads = 1;
    if (threads > n) threads = (int)(n ? n : 1);
    std::vector<int> status(threads, 0);
    std::vector<int64_t> bad(threads, -1);
    auto work = [&](int t) {
        char key[24];
        for (int64_t i = t; i < n; i += threads) {
            const int kl = snprintf(key, sizeof(key), "%lld", (long long)indices[i]);
            const uint8_t* d = nullptr;
            size_t len = 0;
            const int rc = lookup(db, key, (size_t)kl, &d, &len);
            if (rc != 0) { status[t] = rc == 1 ? 1 : 2; bad[t] = indices[i]; return; }
            if (dst) {
                if ((int64_t)len != sizes[i] || offsets[i] < 0 || offsets[i] + (int64_t)len > capacity) { status[t] = 3; bad[t] = indices[i]; return; }
                memcpy((uint8_t*)dst + offsets[i], d, len);
            } else {
                sizes[i] = (int64_t)len;
            }
        }
    };
    if (threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto& th : pool) th.join();
    }
    for (int t = 0; t < threads; ++t) {
        if (status[t] == 1) return tb_fail(TB_ENOENT, "tb_lmdb_gather_ragged: key %lld not found", (long long)bad[t]);
        if (status[t] == 2) return tb_fail(TB_EINVAL, "tb_lmdb_gather_ragged: corrupt page structure at key %lld", (long long)bad[t]);
        if (status[t] == 3) return tb_fail(TB_EINVAL, "tb_lmdb_gather_ragged: record %lld changed size or does not fit the destination", (long long)bad[t]);
    }
    return TB_OK;
}

// ----------------------------------------------------------------------------------------------
// Host helper of T.RandomResizedCrop's parameter replay: the sequential accept / reject walk of
// RandomResizedCrop.get_params (torchvision/transforms/transforms.py) over pre-drawn generator
// words and their per-word candidates.  The candidates (area * uniform(scale), exp(uniform(log
// ratio))) are computed by the caller with torch so that they carry torch's exact float32
// arithmetic; this routine only walks them: two words per attempt, two more for the offsets of an
// accepted box, one for the flip.
//   words, area_cand, aspect_cand : n_words entries each (candidate i belongs to word i)
//   box   : out, int32 [batch][4] = top, left, height, width ; flip : out, uint8 [batch]
//   returns the number of words consumed, or a negative TB_E* code (TB_EINVAL when n_words is too
//   small: the caller draws more and calls again from the start)
// ----------------------------------------------------------------------------------------------
#include <math.h>

// One walk for both layouts: a uniform batch (area_cand[w] = area * u_w, computed by the caller) or a
// ragged one (scale_cand[w] = u_w, multiplied here by every sample's own area).
static int64_t rrc_walk(const uint64_t* words, const double* cand, bool cand_is_area, const double* aspect_cand,
                        int64_t n_words, int64_t batch, const int32_t* heights, const int32_t* widths, int height,
                        int width, double ratio_min, double ratio_max, double flip_p, int32_t* box, uint8_t* flip) {
    int64_t pos = 0;
    for (int64_t b = 0; b < batch; ++b) {
        if (pos + 23 > n_words) return tb_fail(TB_EINVAL, "tb_rrc_params_walk: not enough words");
        const int H = heights ? heights[b] : height, W = widths ? widths[b] : width;
        if (H < 1 || W < 1) return tb_fail(TB_EINVAL, "tb_rrc_params_walk: bad image size");
        const double area = (double)H * (double)W;
        bool found = false;
        for (int attempt = 0; attempt < 10 && !found; ++attempt) {
            const double ta = cand_is_area ? cand[pos] : area * cand[pos], ar = aspect_cand[pos + 1];
            pos += 2;
            // Python's int(round(x)): round half to even
            const long w = lrint(nearbyint(sqrt(ta * ar)));
            const long h = lrint(nearbyint(sqrt(ta / ar)));
            if (w > 0 && w <= W && h > 0 && h <= H) {
                box[4 * b + 0] = (int32_t)(words[pos] % (uint64_t)(H - h + 1));
                box[4 * b + 1] = (int32_t)(words[pos + 1] % (uint64_t)(W - w + 1));
                box[4 * b + 2] = (int32_t)h;
                box[4 * b + 3] = (int32_t)w;
                pos += 2;
                found = true;
            }
        }
        if (!found) {   // central crop fallback
            const double in_ratio = (double)W / (double)H;
            long w, h;
            if (in_ratio < ratio_min) { w = W; h = lrint(nearbyint((double)w / ratio_min)); }
            else if (in_ratio > ratio_max) { h = H; w = lrint(nearbyint((double)h * ratio_max)); }
            else { w = W; h = H; }
            box[4 * b + 0] = (int32_t)((H - h) / 2);
            box[4 * b + 1] = (int32_t)((W - w) / 2);
            box[4 * b + 2] = (int32_t)h;
            box[4 * b + 3] = (int32_t)w;
        }
        flip[b] = 0;
        if (flip_p >= 0.0) {
            const float u = (float)(words[pos] & 0xFFFFFFull) * 5.9604644775390625e-08f;   // 2^-24
            flip[b] = u < (float)flip_p ? 1 : 0;
            pos += 1;
        }
    }
    return pos;
}

extern "C" int64_t tb_rrc_params_walk(const uint64_t* words, const double* area_cand, const double* aspect_cand,
                                      int64_t n_words, int64_t batch, int height, int width, double ratio_min,
                                      double ratio_max, double flip_p, int32_t* box, uint8_t* flip) {
    if (!words || !area_cand || !aspect_cand || !box || !flip || batch < 0 || height < 1 || width < 1)
        return tb_fail(TB_EINVAL, "tb_rrc_params_walk: bad argument");
    return rrc_walk(words, area_cand, true, aspect_cand, n_words, batch, nullptr, nullptr, height, width, ratio_min,
                    ratio_max, flip_p, box, flip);
}

extern "C" int64_t tb_rrc_params_walk_ragged(const uint64_t* words, const double* scale_cand, const double* aspect_cand,
                                             int64_t n_words, int64_t batch, const int32_t* heights,
                                             const int32_t* widths, double ratio_min, double ratio_max, double flip_p,
                                             int32_t* box, uint8_t* flip) {
    if (!words || !scale_cand || !aspect_cand || !box || !flip || batch < 0 || !heights || !widths)
        return tb_fail(TB_EINVAL, "tb_rrc_params_walk_ragged: bad argument");
    return rrc_walk(words, scale_cand, false, aspect_cand, n_words, batch, heights, widths, 0, 0, ratio_min, ratio_max,
                    flip_p, box, flip);
}

// ----------------------------------------------------------------------------------------------
// In-place GPU access to the records: the read-only map is registered with CUDA (page-locked and
// mapped into the device address space), after which kernels can read records straight out of the
// page cache over PCIe -- no gather into a pinned staging buffer at all.
// ----------------------------------------------------------------------------------------------
struct TbLmdbCuda {
    const tb_lmdb* db;
    void* device_base;
};
static TbLmdbCuda g_registered[16];
static std::mutex g_registered_mu;       // register / unregister may come from different threads

extern "C" int tb_lmdb_cuda_register(tb_lmdb* db, void** device_base) {
    TB_REQUIRE(db && device_base, "tb_lmdb_cuda_register: NULL argument");
    std::lock_guard<std::mutex> lock(g_registered_mu);
    for (auto& r : g_registered)
        if (r.db == db) { *device_base = r.device_base; return TB_OK; }
    int ro_supported = 0, dev = 0;
    TB_CUDA(cudaGetDevice(&dev));
    (void)cudaDeviceGetAttribute(&ro_supported, cudaDevAttrHostRegisterReadOnlySupported, dev);
    unsigned flags = cudaHostRegisterMapped | cudaHostRegisterPortable;
    if (ro_supported) flags |= cudaHostRegisterReadOnly;
    // cudaHostRegister refuses file mappings without PROT_WRITE on this platform (cudaErrorInvalidValue even
    // with cudaHostRegisterReadOnly), so the protection of the private mapping is widened here and only here.
    // Nothing ever writes through it.  Cost to know about: registration page-locks the WHOLE file in host
    // memory (and may copy pages into anonymous memory), so this is for databases that fit in RAM.
    if (mprotect((void*)db->map, db->size, PROT_READ | PROT_WRITE) != 0)
        return tb_fail(TB_ENOTSUP, "tb_lmdb_cuda_register: mprotect(%zu bytes, read+write): %s", db->size, strerror(errno));
    cudaError_t e = cudaHostRegister((void*)db->map, db->size, flags);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        mprotect((void*)db->map, db->size, PROT_READ);
        return tb_fail(TB_ENOTSUP, "tb_lmdb_cuda_register: cudaHostRegister(%zu bytes, flags 0x%x) failed: %s "
                       "(read-only registration supported: %d)", db->size, flags, cudaGetErrorString(e), ro_supported);
    }
    void* dptr = nullptr;
    e = cudaHostGetDevicePointer(&dptr, (void*)db->map, 0);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        cudaHostUnregister((void*)db->map);
        mprotect((void*)db->map, db->size, PROT_READ);
        return tb_cuda_fail(e, "cudaHostGetDevicePointer");
    }
    for (auto& r : g_registered)
        if (!r.db) { r.db = db; r.device_base = dptr; *device_base = dptr; return TB_OK; }
    cudaHostUnregister((void*)db->map);
    mprotect((void*)db->map, db->size, PROT_READ);
    return tb_fail(TB_ENOTSUP, "tb_lmdb_cuda_register: too many registered databases");
}

extern "C" int tb_lmdb_cuda_unregister(tb_lmdb* db) {
    std::lock_guard<std::mutex> lock(g_registered_mu);
    for (auto& r : g_registered)
        if (r.db == db) {
            cudaHostUnregister((void*)db->map);
            mprotect((void*)db->map, db->size, PROT_READ);
            r.db = nullptr;
            r.device_base = nullptr;
        }
    return TB_OK;
}

// Byte offsets (from the start of the map) of n fixed-size records `str(idx)`; every record must
// be exactly record_bytes long and start on a 16-byte boundary (true for values LMDB keeps on
// overflow pages: page start + 16).
extern "C" int tb_lmdb_record_offsets(const tb_lmdb* db, const int64_t* indices, int64_t n, int64_t record_bytes,
                                      int64_t* offsets) {
    TB_REQUIRE(db && (n == 0 || (indices && offsets)) && n >= 0, "tb_lmdb_record_offsets: bad argument");
    char key[24];
    for (int64_t i = 0; i < n; ++i) {
        const int kl = snprintf(key, sizeof(key), "%lld", (long long)indices[i]);
        const uint8_t* d = nullptr;
        size_t len = 0;
        const int rc = lookup(db, key, (size_t)kl, &d, &len);
        if (rc == 1) return tb_fail(TB_ENOENT, "tb_lmdb_record_offsets: key %lld not found", (long long)indices[i]);
        if (rc < 0) return tb_fail(TB_EINVAL, "tb_lmdb_record_offsets: corrupt page structure at key %lld", (long long)indices[i]);
        if ((int64_t)len != record_bytes)
            return tb_fail(TB_EINVAL, "tb_lmdb_record_offsets: record %lld is not %lld bytes", (long long)indices[i], (long long)record_bytes);
        offsets[i] = (int64_t)(d - db->map);
    }
    return TB_OK;
}

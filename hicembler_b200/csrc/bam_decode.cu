// Host-side BAM/BGZF -> pair-record decoder.  Replaces the `samtools view` text pipe of
// bam2clusters.py:108-117,139-146 for the device path: BGZF blocks (bin/snap/SNAPLib/Bam.h:322-350) are inflated
// by a pool of zlib workers, alignment records (Bam.h:91-105) are walked once and only the core fields the hot
// path reads are kept -- as 24-byte hc_pair_t or as the 16-byte wire form hc_pair16_t -- straight into a caller
// buffer (pinned memory for the H2D copy).  No filtering here: -q / -F are applied by hc_bin_pairs on the device.
// No device work in this file.
#include <fcntl.h>
#include <stdlib.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <atomic>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

struct Block {
    size_t off;       // offset of the block in the file
    uint32_t csize;   // whole block size (BSIZE + 1)
    uint32_t hdr;     // 12 + XLEN: start of the deflate stream inside the block
    uint32_t isize;   // uncompressed size
};

struct Bam {
    const uint8_t *map = nullptr;
    size_t size = 0;
    int fd = -1;
    std::vector<Block> blocks;
    size_t next_block = 0;
    struct Buf {  // inflated bytes not yet consumed (plain malloc: no zero fill on growth)
        uint8_t *p = nullptr;
        size_t len = 0, cap = 0;
        uint8_t *data() { return p; }
        size_t size() const { return len; }
        bool resize(size_t n) {
            if (n > cap) {
                const size_t c = n + n / 4 + 4096;
                uint8_t *q = (uint8_t *)realloc(p, c);
                if (!q) return false;
                p = q;
                cap = c;
            }
            len = n;
            return true;
        }
        ~Buf() { free(p); }
    } win;
    size_t cur = 0;            // read cursor in win
    int threads = 1;
    bool header_done = false;
    int32_t n_ref = 0;
    int64_t n_records = 0;
    char err[256] = "";
    double t_inflate = 0, t_move = 0;
};

inline uint16_t rd16(const uint8_t *p) { return (uint16_t)(p[0] | (p[1] << 8)); }
inline uint32_t rd32(const uint8_t *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

bool scan_blocks(Bam &b) {
    size_t p = 0;
    while (p < b.size) {
        if (p + 18 > b.size || b.map[p] != 0x1f || b.map[p + 1] != 0x8b || b.map[p + 2] != 8 || !(b.map[p + 3] & 4)) {
            snprintf(b.err, sizeof b.err, "not a BGZF block at offset %zu", p);
            return false;
        }
        const uint32_t xlen = rd16(b.map + p + 10);
        size_t q = p + 12;
        uint32_t bsize = 0;
        while (q + 4 <= p + 12 + xlen && q + 4 <= b.size) {
            const uint32_t slen = rd16(b.map + q + 2);
            if (b.map[q] == 66 && b.map[q + 1] == 67 && slen == 2 && q + 6 <= b.size) bsize = rd16(b.map + q + 4) + 1u;
            q += 4 + slen;
        }
        if (!bsize || p + bsize > b.size || bsize < 12 + xlen + 8) {
            snprintf(b.err, sizeof b.err, "BGZF block without a valid BC subfield at offset %zu", p);
            return false;
        }
        Block k;
        k.off = p;
        k.csize = bsize;
        k.hdr = 12 + xlen;
        k.isize = rd32(b.map + p + bsize - 4);
        b.blocks.push_back(k);
        p += bsize;
    }
    return true;
}

// inflate blocks [b0, b1) to dst (consecutive), in parallel; returns false on a corrupt block
bool inflate_range(Bam &b, size_t b0, size_t b1, uint8_t *dst) {
    std::vector<size_t> uoff(b1 - b0 + 1, 0);
    for (size_t k = b0; k < b1; ++k) uoff[k - b0 + 1] = uoff[k - b0] + b.blocks[k].isize;
    std::atomic<size_t> next(b0);
    std::atomic<int> bad(0);
    auto work = [&]() {
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= b1 || bad.load()) return;
            const Block &blk = b.blocks[k];
            if (blk.isize == 0) continue;
            z_stream zs;
            memset(&zs, 0, sizeof zs);
            if (inflateInit2(&zs, -15) != Z_OK) { bad = 1; return; }
            zs.next_in = const_cast<Bytef *>(b.map + blk.off + blk.hdr);
            zs.avail_in = blk.csize - blk.hdr - 8;
            zs.next_out = dst + uoff[k - b0];
            zs.avail_out = blk.isize;
            const int rc = inflate(&zs, Z_FINISH);
            inflateEnd(&zs);
            const uint32_t crc = rd32(b.map + blk.off + blk.csize - 8);
            if (rc != Z_STREAM_END || zs.total_out != blk.isize ||
                (uint32_t)crc32(crc32(0L, Z_NULL, 0), dst + uoff[k - b0], blk.isize) != crc)
                bad = 1;
        }
    };
    const int nt = (int)std::min<size_t>((size_t)std::max(1, b.threads), b1 - b0);
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(work);
    work();
    for (auto &t : pool) t.join();
    if (bad.load()) snprintf(b.err, sizeof b.err, "corrupt BGZF block (inflate / CRC / size) in blocks %zu..%zu", b0, b1);
    return !bad.load();
}

constexpr size_t BATCH_BLOCKS = 2048;  // <= 128 MB of inflated data per refill

// make at least `need` unread bytes available; false at end of file (or on error: b.err set)
bool ensure(Bam &b, size_t need) {
    while (b.win.size() - b.cur < need) {
        if (b.next_block >= b.blocks.size()) return false;
        const double t0 = hc::wall_seconds();
        if (b.cur > 0) {  // drop what was consumed
            memmove(b.win.data(), b.win.data() + b.cur, b.win.size() - b.cur);
            b.win.resize(b.win.size() - b.cur);
            b.cur = 0;
        }
        const size_t b0 = b.next_block, b1 = std::min(b.blocks.size(), b0 + BATCH_BLOCKS);
        size_t add = 0;
        for (size_t k = b0; k < b1; ++k) add += b.blocks[k].isize;
        const size_t old = b.win.size();
        if (!b.win.resize(old + add)) { snprintf(b.err, sizeof b.err, "out of memory"); return false; }
        const double t1 = hc::wall_seconds();
        if (!inflate_range(b, b0, b1, b.win.data() + old)) return false;
        b.t_inflate += hc::wall_seconds() - t1;
        b.t_move += t1 - t0;
        b.next_block = b1;
    }
    return true;
}

bool read_header(Bam &b) {
    if (!ensure(b, 12)) { if (!b.err[0]) snprintf(b.err, sizeof b.err, "truncated BAM header"); return false; }
    if (memcmp(b.win.data() + b.cur, "BAM\1", 4) != 0) { snprintf(b.err, sizeof b.err, "bad BAM magic"); return false; }
    const uint32_t l_text = rd32(b.win.data() + b.cur + 4);
    if (!ensure(b, 12 + (size_t)l_text)) { if (!b.err[0]) snprintf(b.err, sizeof b.err, "truncated BAM header text"); return false; }
    b.n_ref = (int32_t)rd32(b.win.data() + b.cur + 8 + l_text);
    b.cur += 12 + (size_t)l_text;
    for (int32_t r = 0; r < b.n_ref; ++r) {
        if (!ensure(b, 4)) { if (!b.err[0]) snprintf(b.err, sizeof b.err, "truncated reference list"); return false; }
        const uint32_t l_name = rd32(b.win.data() + b.cur);
        if (!ensure(b, 8 + (size_t)l_name)) { if (!b.err[0]) snprintf(b.err, sizeof b.err, "truncated reference list"); return false; }
        b.cur += 8 + (size_t)l_name;
    }
    b.header_done = true;
    return true;
}

}  // namespace

extern "C" int hc_bam_open(const char *path, int32_t n_threads, void **h_handle, int32_t *h_n_ref,
                           int64_t *h_uncompressed_bytes) {
    HC_REQUIRE(path && h_handle, "bad arguments");
    Bam *b = new Bam();
    b->threads = n_threads > 0 ? n_threads : (int)std::max(1u, std::thread::hardware_concurrency());
    b->fd = open(path, O_RDONLY);
    struct stat st;
    if (b->fd < 0 || fstat(b->fd, &st) != 0) {
        hc::set_error("hc_bam_open: cannot open %s", path);
        if (b->fd >= 0) close(b->fd);
        delete b;
        return HC_EINVAL;
    }
    b->size = (size_t)st.st_size;
    if (b->size) {
        void *m = mmap(nullptr, b->size, PROT_READ, MAP_PRIVATE, b->fd, 0);
        if (m == MAP_FAILED) {
            hc::set_error("hc_bam_open: mmap failed for %s", path);
            close(b->fd);
            delete b;
            return HC_EINVAL;
        }
        b->map = (const uint8_t *)m;
        madvise(m, b->size, MADV_SEQUENTIAL);
    }
    if (!scan_blocks(*b) || !read_header(*b)) {
        hc::set_error("hc_bam_open: %s: %s", path, b->err[0] ? b->err : "empty file");
        if (b->map) munmap((void *)b->map, b->size);
        close(b->fd);
        delete b;
        return HC_EINVAL;
    }
    if (h_n_ref) *h_n_ref = b->n_ref;
    if (h_uncompressed_bytes) {
        int64_t t = 0;
        for (const Block &k : b->blocks) t += k.isize;
        *h_uncompressed_bytes = t;
    }
    *h_handle = b;
    return HC_OK;
}

// Up to `capacity` further records into h_out (wire 24: hc_pair_t, wire 16: hc_pair16_t) and, if not NULL, TLEN
// and the CIGAR reference span (>= 1) into h_tlen / h_ref_span.  *h_n = records written; 0 = end of file.
extern "C" int hc_bam_next(void *handle, int32_t wire, void *h_out, int64_t capacity, int32_t *h_tlen,
                           int32_t *h_ref_span, int64_t *h_n) {
    HC_REQUIRE(handle && h_n && (wire == 16 || wire == 24) && capacity >= 0 && (capacity == 0 || h_out), "bad arguments");
    Bam &b = *(Bam *)handle;
    int64_t n = 0;
    hc_pair_t *o24 = (hc_pair_t *)h_out;
    hc_pair16_t *o16 = (hc_pair16_t *)h_out;
    while (n < capacity) {
        if (!ensure(b, 4)) break;
        const uint32_t bs = rd32(b.win.data() + b.cur);
        if (bs < 32 || !ensure(b, 4 + (size_t)bs)) {
            if (!b.err[0]) snprintf(b.err, sizeof b.err, "truncated alignment record %lld", (long long)b.n_records);
            hc::set_error("hc_bam_next: %s", b.err);
            return HC_EINVAL;
        }
        const uint8_t *p = b.win.data() + b.cur + 4;  // core, Bam.h:91-105
        const int32_t refID = (int32_t)rd32(p), pos = (int32_t)rd32(p + 4);
        const uint32_t l_read_name = p[8], mapq = p[9], n_cigar = rd16(p + 12), flag = rd16(p + 14);
        const int32_t l_seq = (int32_t)rd32(p + 16), nref = (int32_t)rd32(p + 20), npos = (int32_t)rd32(p + 24);
        const int32_t tlen = (int32_t)rd32(p + 28);
        if (32 + (size_t)l_read_name + 4 * (size_t)n_cigar > bs) {
            hc::set_error("hc_bam_next: record %lld: CIGAR beyond the record", (long long)b.n_records);
            return HC_EINVAL;
        }
        if (wire == 24) {
            hc_pair_t r;
            r.refID = refID; r.pos = pos; r.next_refID = nref; r.next_pos = npos; r.l_seq = l_seq;
            r.flag = (uint16_t)flag; r.mapq = (uint8_t)mapq; r.pad = 0;
            o24[n] = r;
        } else {
            if (refID >= 0xfffff || nref >= 0xfffff) {
                hc::set_error("hc_bam_next: reference id >= 2^20 - 1 does not fit hc_pair16_t, use wire 24");
                return HC_EINVAL;
            }
            const int32_t half = (l_seq > 1 ? l_seq : 1) >> 1;
            hc_pair16_t r;
            r.mid = (uint32_t)(pos + half);
            r.next_mid = (uint32_t)(npos + half);
            r.ref_flag = ((uint32_t)refID & 0xfffffu) | ((flag & 0xfffu) << 20);
            r.nref_mapq = ((uint32_t)nref & 0xfffffu) | (mapq << 20);
            o16[n] = r;
        }
        if (h_tlen) h_tlen[n] = tlen;
        if (h_ref_span) {
            const uint8_t *c = p + 32 + l_read_name;
            int64_t span = 0;
            for (uint32_t k = 0; k < n_cigar; ++k) {
                const uint32_t v = rd32(c + 4 * k), op = v & 15u;
                if (op == 0 || op == 2 || op == 3 || op == 7 || op == 8) span += v >> 4;  // M D N = X
            }
            h_ref_span[n] = span ? (int32_t)span : 1;
        }
        b.cur += 4 + (size_t)bs;
        b.n_records++;
        n++;
    }
    if (b.err[0]) {
        hc::set_error("hc_bam_next: %s", b.err);
        return HC_EINVAL;
    }
    if (n < capacity && b.next_block >= b.blocks.size() && b.win.size() - b.cur != 0) {
        hc::set_error("hc_bam_next: %zu stray bytes after the last alignment record", b.win.size() - b.cur);
        return HC_EINVAL;
    }
    *h_n = n;
    return HC_OK;
}

extern "C" int hc_bam_close(void *handle) {
    if (!handle) return HC_OK;
    Bam *b = (Bam *)handle;
    if (getenv("HC_BAM_TRACE"))
        fprintf(stderr, "[bam trace] %lld records, inflate %.3f s (%d threads), buffer moves %.3f s\n",
                (long long)b->n_records, b->t_inflate, b->threads, b->t_move);
    if (b->map) munmap((void *)b->map, b->size);
    if (b->fd >= 0) close(b->fd);
    delete b;
    return HC_OK;
}

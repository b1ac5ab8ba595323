// coral_b200/csrc/bam_reader.cpp — host-side BGZF/BAM ingest (replaces pysam.AlignmentFile(path,'rb'),
// reference infer_breakpoint_graph.py:65).  I/O only: inflates BGZF members in parallel with zlib and
// frames the alignment records into one (pinned when CUDA is present) host buffer with 16-byte
// aligned record starts.  No htslib exists in this image; format per SAM/BAM spec.
#include <cuda_runtime.h>
#include <zlib.h>

#include <atomic>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/coral_b200.h"

struct coral_bam {
    std::vector<std::string> ref_names;
    std::vector<int64_t> ref_lens;
    std::string header_text;
    uint8_t* recs = nullptr;
    uint64_t nbytes = 0;
    bool pinned = false;
    std::vector<uint64_t> offsets;
};

namespace {

struct Member { uint64_t in_off, in_len, out_off, out_len; };

bool inflate_member(const uint8_t* src, uint64_t len, uint8_t* dst, uint64_t dlen) {
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    if (inflateInit2(&zs, -15) != Z_OK) return false;
    zs.next_in = const_cast<Bytef*>(src);
    zs.avail_in = (uInt)len;
    zs.next_out = dst;
    zs.avail_out = (uInt)dlen;
    int rc = inflate(&zs, Z_FINISH);
    inflateEnd(&zs);
    return rc == Z_STREAM_END && zs.total_out == dlen;
}

inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
inline uint16_t rd16(const uint8_t* p) { uint16_t v; memcpy(&v, p, 2); return v; }

void set_err(char* err, size_t n, const std::string& s) {
    if (err && n) { snprintf(err, n, "%s", s.c_str()); }
}

}  // namespace

extern "C" {

int coral_bam_read(const char* path, int n_threads, coral_bam** out, char* err, size_t err_len) {
    if (!path || !out) return CORAL_E_ARG;
    *out = nullptr;
    FILE* fp = fopen(path, "rb");
    if (!fp) { set_err(err, err_len, std::string("cannot open ") + path); return CORAL_E_IO; }
    fseek(fp, 0, SEEK_END);
    const uint64_t fsz = (uint64_t)ftell(fp);
    fseek(fp, 0, SEEK_SET);
    std::vector<uint8_t> raw(fsz);
    if (fsz && fread(raw.data(), 1, fsz, fp) != fsz) { fclose(fp); set_err(err, err_len, "short read"); return CORAL_E_IO; }
    fclose(fp);

    // 1. member table
    std::vector<Member> mem;
    uint64_t p = 0, total = 0;
    while (p < fsz) {
        if (p + 18 > fsz || raw[p] != 0x1f || raw[p + 1] != 0x8b || raw[p + 2] != 8 || !(raw[p + 3] & 4)) {
            set_err(err, err_len, "not a BGZF member at offset " + std::to_string(p));
            return CORAL_E_FORMAT;
        }
        const uint32_t xlen = rd16(&raw[p + 10]);
        uint64_t q = p + 12, xend = p + 12 + xlen;
        int64_t bsize = -1;
        while (q + 4 <= xend && xend <= fsz) {
            const uint32_t slen = rd16(&raw[q + 2]);
            if (q + 4 + slen > xend) break;                       // a subfield that overruns the extra field: not read
            if (raw[q] == 'B' && raw[q + 1] == 'C' && slen == 2) bsize = rd16(&raw[q + 4]);
            q += 4 + slen;
        }
        if (bsize < 0 || p + (uint64_t)bsize + 1 > fsz) { set_err(err, err_len, "bad BGZF BC field"); return CORAL_E_FORMAT; }
        const uint64_t mend = p + (uint64_t)bsize + 1;
        // the member must hold its header, extra field, CRC32 and ISIZE: otherwise the CDATA length below underflows
        if (xend > fsz || mend < xend + 8) { set_err(err, err_len, "BGZF member shorter than its own header at offset " + std::to_string(p)); return CORAL_E_FORMAT; }
        const uint32_t isize = rd32(&raw[mend - 4]);
        mem.push_back({xend, mend - 8 - xend, total, isize});
        total += isize;
        p = mend;
    }
    // 2. parallel inflate
    std::vector<uint8_t> data(total + 8);
    int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    std::atomic<size_t> next{0};
    std::atomic<int> failed{0};
    auto work = [&]() {
        for (;;) {
            size_t i = next.fetch_add(16);
            if (i >= mem.size()) break;
            for (size_t j = i; j < std::min(i + 16, mem.size()); j++) {
                const Member& m = mem[j];
                if (m.out_len && !inflate_member(&raw[m.in_off], m.in_len, &data[m.out_off], m.out_len)) failed = 1;
            }
        }
    };
    {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; t++) th.emplace_back(work);
        for (auto& x : th) x.join();
    }
    if (failed) { set_err(err, err_len, "inflate failed"); return CORAL_E_FORMAT; }
    raw.clear();
    raw.shrink_to_fit();

    // 3. header
    coral_bam* b = new coral_bam();
    if (total < 12 || memcmp(data.data(), "BAM\1", 4) != 0) { delete b; set_err(err, err_len, "not a BAM stream"); return CORAL_E_FORMAT; }
    uint64_t o = 4;
    const uint32_t l_text = rd32(&data[o]); o += 4;
    if (o + l_text + 4 > total) { delete b; set_err(err, err_len, "truncated header"); return CORAL_E_FORMAT; }
    b->header_text.assign((const char*)&data[o], l_text); o += l_text;
    const uint32_t n_ref = rd32(&data[o]); o += 4;
    for (uint32_t i = 0; i < n_ref; i++) {
        if (o + 4 > total) { delete b; set_err(err, err_len, "truncated reference table"); return CORAL_E_FORMAT; }
        const uint32_t ln = rd32(&data[o]); o += 4;
        if (o + ln + 4 > total || ln == 0) { delete b; set_err(err, err_len, "truncated reference table"); return CORAL_E_FORMAT; }
        b->ref_names.emplace_back((const char*)&data[o], ln - 1); o += ln;
        b->ref_lens.push_back((int32_t)rd32(&data[o])); o += 4;
    }
    // 4. frame records
    std::vector<uint64_t> src;
    uint64_t dst = 0;
    while (o + 4 <= total) {
        const uint32_t bs = rd32(&data[o]);
        if (bs < 32 || o + 4 + bs > total) { delete b; set_err(err, err_len, "truncated alignment record"); return CORAL_E_FORMAT; }
        src.push_back(o);
        b->offsets.push_back(dst);
        dst += ((uint64_t)bs + 4 + 15) & ~uint64_t(15);
        o += 4 + (uint64_t)bs;
    }
    b->offsets.push_back(dst);
    b->nbytes = dst;
    const uint64_t alloc = dst + 64;
    void* hp = nullptr;
    if (cudaHostAlloc(&hp, alloc, cudaHostAllocDefault) == cudaSuccess) {
        b->pinned = true;
    } else {
        cudaGetLastError();  // clear; plain memory when no CUDA device is present (CPU-only tooling)
        hp = aligned_alloc(64, (alloc + 63) & ~uint64_t(63));
        if (!hp) { delete b; set_err(err, err_len, "out of host memory"); return CORAL_E_NOMEM; }
    }
    b->recs = (uint8_t*)hp;
    const size_t nrec = src.size();
    std::atomic<size_t> nx{0};
    auto cp = [&]() {
        for (;;) {
            size_t i = nx.fetch_add(256);
            if (i >= nrec) break;
            for (size_t j = i; j < std::min(i + 256, nrec); j++) {
                const uint64_t len = 4 + (uint64_t)rd32(&data[src[j]]);
                memcpy(b->recs + b->offsets[j], &data[src[j]], len);
                const uint64_t padn = b->offsets[j + 1] - b->offsets[j] - len;
                if (padn) memset(b->recs + b->offsets[j] + len, 0, padn);
            }
        }
    };
    {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; t++) th.emplace_back(cp);
        for (auto& x : th) x.join();
    }
    memset(b->recs + dst, 0, 64);
    *out = b;
    return CORAL_OK;
}

void coral_bam_free(coral_bam* b) {
    if (!b) return;
    if (b->recs) { if (b->pinned) cudaFreeHost(b->recs); else free(b->recs); }
    delete b;
}
int32_t coral_bam_n_ref(const coral_bam* b) { return b ? (int32_t)b->ref_names.size() : -1; }
const char* coral_bam_ref_name(const coral_bam* b, int32_t tid) {
    return (b && tid >= 0 && (size_t)tid < b->ref_names.size()) ? b->ref_names[tid].c_str() : nullptr;
}
int64_t coral_bam_ref_len(const coral_bam* b, int32_t tid) {
    return (b && tid >= 0 && (size_t)tid < b->ref_lens.size()) ? b->ref_lens[tid] : -1;
}
int64_t coral_bam_n_records(const coral_bam* b) { return b ? (int64_t)b->offsets.size() - 1 : -1; }
const uint8_t* coral_bam_records(const coral_bam* b, uint64_t* nbytes) {
    if (!b) return nullptr;
    if (nbytes) *nbytes = b->nbytes;
    return b->recs;
}
const uint64_t* coral_bam_offsets(const coral_bam* b) { return b ? b->offsets.data() : nullptr; }

}  // extern "C"

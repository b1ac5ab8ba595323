// coral_b200/csrc/tables.h — POD rows shared between kernels and the ABI plumbing.
#pragma once
#include <cstdint>

namespace coral {

// One parsed SA:Z entry `rname,pos,strand,CIGAR,mapQ,NM` (cigar_parsing.py:247).
struct SaEntry {
    uint64_t h1, h2;   // hash of the raw entry bytes (exact-string de-duplication)
    int64_t pos1;      // 1-based position as written
    int64_t num[4];    // CIGAR numbers in order
    int64_t nm;        // last field
    uint32_t rec;      // record the entry was read from
    int32_t tid;       // reference id of rname, -1 if not in the header
    int32_t mapq;
    uint8_t strand;    // 0 '+', 1 anything else
    uint8_t sig;       // 0..8 signature index; 100 reject, 101 KeyError, 102 malformed
    uint8_t pad[2];
};

// read status bits (coral_chimeric_reads)
enum : int32_t {
    READ_OK = 0,
    READ_NO_PRIMARY = 1,   // ibg:164-169: dropped with a warning
    READ_REJECTED = 2,     // cp:248-253: entry without S or M => ([], [], [])
    READ_KEYERROR = 3,     // cp:255: signature outside the nine keys
    READ_MALFORMED = 4,    // entry text the reference would fail to int()/index
};

}  // namespace coral

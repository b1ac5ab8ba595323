// coral_b200/csrc/msg.cuh — wire layout of the packed exchange messages (exchange.cu, stab.cu).
#pragma once
#include <cstdint>

namespace coral {

struct MsgHdr {
    uint64_t n_rows, text_bytes;
    double nm_sum, nm_sq;
    int64_t nm_n, nm_zd;
    uint64_t kind, pad;
};
static_assert(sizeof(MsgHdr) == 64, "MsgHdr is 64 bytes on the wire");

struct RowMsg {               // one record that matters to fetch: SA-bearing, or first primary of a chimeric name
    uint64_t lo, hi;          // 128-bit read-name hash
    uint32_t flag_mapq, l_seq;
    int32_t nm;
    uint32_t rec;             // record index inside the sender's shard (orders the two messages of a rank)
    uint32_t name_len, sa_len;
    uint64_t text_off;        // of the row's slot (name incl. NUL, then SA text) inside the message text
};
static_assert(sizeof(RowMsg) == 48, "RowMsg is 48 bytes on the wire");

struct NameRow { uint64_t edge, lo, hi, mask; };   // concordant exchange: one (edge, read, covered points) row

enum { MSG_SA = 1, MSG_EXTRA = 2, MSG_NAMES = 3 };

}  // namespace coral

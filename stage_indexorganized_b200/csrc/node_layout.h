// node_layout.h — byte layout of Stage's BzTree nodes as the GPU path consumes it.
//
// Restated from the reference (paths relative to the reference repository):
//   BaseNode {vptr, bool is_leaf, NodeHeader header, RecordMetadata record_metadata[]}
//                                                        include/vstore/b_tree.h:109-113
//   NodeHeader {u32 size, u32 sorted_count, atomic<u32> next_record_slot, StatusWord status}
//                                                        include/vstore/version_store.h:158-231
//   RecordMetadata {u64 meta, u64 next_ptr, RecordLocation *loc_ptr}
//                                                        include/vstore/record_meta.h:30-70
//   TupleHeader (56 B)                                   include/vstore/version_store.h:28-46
// Offsets/sizes were verified by compiling the reference headers.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define STAGE_HD __host__ __device__ __forceinline__
#else
#define STAGE_HD inline
#endif

namespace stage {

enum : uint32_t {
    kOffIsLeaf = 8,
    kOffNodeSize = 16,
    kOffSortedCount = 20,
    kOffStatus = 32,
    kOffMeta = 40,      // record_metadata[0]
    kMetaSize = 24,     // sizeof(RecordMetadata)
    kMetaNextPtr = 8,   // RecordMetadata.next_ptr
    kTupleHeaderSize = 56,
    kThBegin = 4,       // TupleHeader.begin_id_
    kThCommit = 8,      // TupleHeader.comm_id_
    kThNext = 16,       // TupleHeader.next_
    kThSlot = 40        // TupleHeader.tuple_slot_
};

// RecordMetadata.meta bit fields (record_meta.h:66-70). NB the key-length mask really is 0x2FFF.
constexpr uint64_t kMetaControl = 1ull << 63;
constexpr uint64_t kMetaVisible = 1ull << 62;
constexpr uint64_t kMetaKeyLen = 0x2FFFull << 48;
constexpr uint64_t kMetaOffset = 0xFFFFull << 32;
constexpr uint64_t kMetaTxn = 0xFFFFFFFFull;
// NodeHeader::StatusWord (version_store.h:176-180)
constexpr uint64_t kStatusFrozen = 1ull << 60;

STAGE_HD uint32_t meta_key_len(uint64_t m) { return (uint32_t) ((m & kMetaKeyLen) >> 48); }
STAGE_HD uint32_t pad_key_len(uint32_t k) { return (k + 7u) / 8u * 8u; }   // record_meta.h:89-91
STAGE_HD uint32_t meta_offset(uint64_t m) { return (uint32_t) ((m & kMetaOffset) >> 32); }
STAGE_HD bool meta_visible(uint64_t m) { return (m & kMetaVisible) != 0; }
STAGE_HD bool meta_inflight(uint64_t m) { return (m & kMetaVisible) && (m & kMetaControl); }
STAGE_HD uint32_t meta_commit_id(uint64_t m) { return (uint32_t) (m & kMetaTxn); }
STAGE_HD uint32_t status_record_count(uint64_t s) { return (uint32_t) ((s >> 44) & 0xFFFF); }

}  // namespace stage

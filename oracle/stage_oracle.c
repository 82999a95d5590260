/* oracle/stage_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE (see stage_oracle.h).
 *
 * CPU restatement, in plain C over raw node bytes, of the reference's index-organized read path:
 * key order, root-to-leaf routing, in-leaf search, leaf range scan, the leaf-hopping iterator and the
 * MVCC visibility rule — plus a single-threaded restatement of the loader (leaf insert, leaf split,
 * inner-node growth/split) so that byte-identical trees can be built where /root/reference is absent.
 * All citations are relative to /root/reference.
 *
 * PARITY PINNED (tests/test_oracle_golden.py): against golden vectors produced by the live, unmodified reference
 * (tests/golden/make_golden.py; MVCC range scans there run the reference's own IndexScanExecutor), against the
 * reference's known-answer tests for this path (test/testing_btree.cpp), and against the live reference library
 * where it is present (leaf images byte for byte, probes, scans, read ids straddling the commit ids, deletes).
 *
 * Node byte layout restated from include/vstore/b_tree.h:109-113, version_store.h:158-231,
 * record_meta.h:30-70 (sizes verified by compiling the reference):
 *   node+8  bool is_leaf        node+16 u32 size        node+20 u32 sorted_count
 *   node+32 u64 status          node+40 RecordMetadata[] (24 B: u64 meta, u64 next_ptr, u64 loc_ptr)
 *   status: bit60 frozen, 44..59 record count, 22..43 block size, 0..21 deleted size
 *   meta  : bit63 control(in-flight), bit62 visible, key length mask 0x2FFF<<48, 32..47 offset,
 *           0..31 commit id
 */
#define _GNU_SOURCE
#include "stage_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <sched.h>
#include <time.h>

/* ------------------------------------------------------------------------------------------ */
/* layout helpers                                                                              */
/* ------------------------------------------------------------------------------------------ */
enum { OFF_IS_LEAF = 8, OFF_SIZE = 16, OFF_SORTED = 20, OFF_STATUS = 32, OFF_META = 40, META_SZ = 24,
       LEAF_HDR = 40, INNER_HDR = 48 };

#define M_CONTROL (1ull << 63)
#define M_VISIBLE (1ull << 62)
#define M_KEYLEN  (0x2FFFull << 48)          /* record_meta.h:68 — bit 12 of the length is dropped */
#define M_OFFSET  (0xFFFFull << 32)
#define M_TXN     0xFFFFFFFFull
#define S_FROZEN  (1ull << 60)
#define S_COUNT   (0xFFFFull << 44)
#define S_BLOCK   (0x3FFFFFull << 22)

typedef uint8_t u8;

static inline uint64_t *meta_at(const u8 *node, uint32_t i) { return (uint64_t *) (node + OFF_META + (size_t) i * META_SZ); }
static inline uint32_t m_keylen(uint64_t m) { return (uint32_t) ((m & M_KEYLEN) >> 48); }
static inline uint32_t pad8(uint32_t k) { return (k + 7u) / 8u * 8u; }            /* record_meta.h:89-91 */
static inline uint32_t m_padlen(uint64_t m) { return pad8(m_keylen(m)); }
static inline uint32_t m_offset(uint64_t m) { return (uint32_t) ((m & M_OFFSET) >> 32); }
static inline int m_visible(uint64_t m) { return (m & M_VISIBLE) != 0; }
static inline int m_inserting(uint64_t m) { return m_visible(m) && (m & M_CONTROL) != 0; } /* record_meta.h:178-181 */
static inline uint32_t m_cid(uint64_t m) { return (uint32_t) (m & M_TXN); }
static inline uint64_t m_finalize(uint64_t off, uint64_t klen, uint64_t cid) {        /* record_meta.h:128-137 */
    uint64_t m = (klen << 48) | (1ull << 62) | (off << 32) | cid;
    return m & ~M_CONTROL;
}
static inline int n_is_leaf(const u8 *n) { return n[OFF_IS_LEAF] != 0; }
static inline uint32_t n_size(const u8 *n) { uint32_t v; memcpy(&v, n + OFF_SIZE, 4); return v; }
static inline uint32_t n_sorted(const u8 *n) { uint32_t v; memcpy(&v, n + OFF_SORTED, 4); return v; }
static inline uint64_t n_status(const u8 *n) { uint64_t v; memcpy(&v, n + OFF_STATUS, 8); return v; }
static inline void n_set_status(u8 *n, uint64_t s) { memcpy(n + OFF_STATUS, &s, 8); }
static inline uint32_t s_count(uint64_t s) { return (uint32_t) ((s & S_COUNT) >> 44) & 0xFFFF; }
static inline uint32_t s_block(uint64_t s) { return (uint32_t) ((s & S_BLOCK) >> 22); }

struct so_tree {
    u8 *root;
    uint32_t payload_size, split_threshold, merge_threshold, leaf_node_size;
    uint64_t n_leaves;
    void **chain_allocs;
    size_t n_chain, cap_chain;
};

/* ------------------------------------------------------------------------------------------ */
/* a1: key order — BaseNode::KeyCompare / my_memcmp (include/vstore/b_tree.h:99-134)           */
/* ------------------------------------------------------------------------------------------ */
int so_key_compare(const void *k1, uint32_t s1, const void *k2, uint32_t s2) {
    if (!k1) return -1;
    if (!k2) return 1;
    uint32_t m = s1 < s2 ? s1 : s2;
    int cmp = 0;
    if (m < 16) {                                  /* signed-char bytewise difference, byte 0 first */
        const signed char *a = (const signed char *) k1, *b = (const signed char *) k2;
        for (uint32_t i = 0; i < m; i++)
            if (a[i] != b[i]) { cmp = a[i] - b[i]; break; }
    } else {
        cmp = memcmp(k1, k2, m);
    }
    if (cmp == 0) return (int) (s1 - s2);
    return cmp;
}

/* key pointer of a record: BaseNode::GetKey (b_tree.h:188-193) / GetRawRecord (b_tree.h:372-392) */
static inline const u8 *leaf_key(const u8 *node, uint64_t m) { return m_visible(m) ? node + m_offset(m) : NULL; }
static inline const u8 *inner_key(const u8 *node, uint64_t m) { return m_padlen(m) == 0 ? NULL : node + m_offset(m); }
static inline u8 *inner_child(const u8 *node, uint64_t m) {
    uint64_t p; memcpy(&p, node + m_offset(m) + m_padlen(m), 8); return (u8 *) (uintptr_t) p;
}

/* ------------------------------------------------------------------------------------------ */
/* a3: InternalNode::GetChildIndex (src/vstore/b_tree.cpp:664-702)                              */
/* ------------------------------------------------------------------------------------------ */
static uint32_t get_child_index(const u8 *node, const void *key, uint32_t ks, int get_le) {
    int32_t left = 0, right = (int32_t) n_sorted(node) - 1, mid;
    for (;;) {
        mid = (left + right) / 2;
        uint64_t m = *meta_at(node, (uint32_t) mid);
        int cmp = so_key_compare(key, ks, inner_key(node, m), m_keylen(m));
        if (cmp == 0) return (uint32_t) (get_le ? mid - 1 : mid);
        if (left > right) return (uint32_t) ((cmp <= 0 && get_le) ? mid - 1 : mid);
        if (cmp > 0) left = mid + 1; else right = mid - 1;
    }
}

typedef struct { u8 *node; uint32_t meta_index; } frame_t;
typedef struct { frame_t f[32]; uint32_t n; u8 *root; } stack_t_;

/* a2: BTree::TraverseToLeaf (src/vstore/b_tree.cpp:1804-1846) */
static u8 *traverse_to_leaf(so_tree *t, stack_t_ *st, const void *key, uint32_t ks, int le_child) {
    u8 *node = t->root;
    if (st) { st->root = node; }
    while (!n_is_leaf(node)) {
        uint32_t idx = get_child_index(node, key, ks, le_child);
        u8 *child = inner_child(node, *meta_at(node, idx));
        if (!child) return NULL;
        if (st) { st->f[st->n].node = node; st->f[st->n].meta_index = idx; st->n++; }
        node = child;
    }
    return node;
}

/* ------------------------------------------------------------------------------------------ */
/* a4: BaseNode::SearchRecordMeta (src/vstore/b_tree.cpp:18-122) with check_concurrency = true  */
/* returns slot index or -1                                                                    */
/* ------------------------------------------------------------------------------------------ */
static int64_t search_record_meta(const u8 *node, const void *key, uint32_t ks) {
    uint32_t sorted = n_sorted(node);
    for (uint32_t i = 0; i < sorted; i++) {
        uint64_t m = *meta_at(node, i);
        if (m == 0) continue;
        if (so_key_compare(key, ks, leaf_key(node, m), m_keylen(m)) == 0 && m_visible(m)) return i;
    }
    uint32_t count = s_count(n_status(node));
    for (uint32_t i = sorted; i < count; i++) {
        uint64_t m = *meta_at(node, i);
        if (!m_visible(m)) continue;
        uint32_t len = m_keylen(m);
        if (len == ks && so_key_compare(key, ks, leaf_key(node, m), len) == 0) return i;
    }
    return -1;
}

/* ------------------------------------------------------------------------------------------ */
/* loader restatement (write path; used only to build tables for the checker)                  */
/* ------------------------------------------------------------------------------------------ */
static u8 *leaf_new(so_tree *t, uint32_t size) {             /* LeafNode::New, b_tree.cpp:791-797 */
    size_t alloc = size < 65536 ? 65536 : size;              /* DramBlock is 64 KB, valen_buffer.h:312 */
    u8 *n = (u8 *) aligned_alloc(4096, alloc);
    memset(n, 0, alloc);
    n[OFF_IS_LEAF] = 1;
    memcpy(n + OFF_SIZE, &size, 4);
    t->n_leaves++;
    return n;
}

so_tree *so_tree_create(uint32_t payload_size, uint32_t split_threshold, uint32_t merge_threshold,
                        uint32_t leaf_node_size) {
    so_tree *t = (so_tree *) calloc(1, sizeof(*t));
    t->payload_size = payload_size;
    t->split_threshold = split_threshold;
    t->merge_threshold = merge_threshold;
    t->leaf_node_size = leaf_node_size;
    t->root = leaf_new(t, leaf_node_size);                   /* b_tree.h:793-803 */
    return t;
}

static void free_subtree(u8 *n) {
    if (!n) return;
    if (!n_is_leaf(n)) {
        uint32_t c = n_sorted(n);
        for (uint32_t i = 0; i < c; i++) free_subtree(inner_child(n, *meta_at(n, i)));
    }
    free(n);
}
void so_tree_destroy(so_tree *t) {
    if (!t) return;
    free_subtree(t->root);
    for (size_t i = 0; i < t->n_chain; i++) free(t->chain_allocs[i]);
    free(t->chain_allocs);
    free(t);
}
const void *so_root(so_tree *t) { return t->root; }
uint64_t so_leaf_count(so_tree *t) { return t->n_leaves; }

enum { RC_OK = 1, RC_KEY_EXISTS = 2, RC_NOT_FOUND = 3, RC_FROZEN = 4, RC_NO_SPACE = 6 };

/* LeafNode::Insert (b_tree.cpp:809-947), single-threaded: the CAS pairs always succeed */
static int leaf_insert(so_tree *t, u8 *leaf, const void *key, uint32_t ks, const void *payload, uint32_t cid,
                       uint64_t **meta_out) {
    uint64_t status = n_status(leaf);
    if (status & S_FROZEN) return RC_FROZEN;
    int64_t dup = search_record_meta(leaf, key, ks);         /* CheckUnique, b_tree.cpp:1395-1417 */
    if (dup >= 0) return RC_KEY_EXISTS;
    uint32_t P = t->payload_size, padk = pad8(ks);
    uint32_t used = LEAF_HDR + s_block(status) + s_count(status) * META_SZ;     /* GetUsedSpace b_tree.h:577-582 */
    uint32_t new_size = used + META_SZ + padk + P;
    if (new_size >= t->leaf_node_size) return RC_NO_SPACE;   /* threshold = leaf_node_size, b_tree.cpp:1869-1871 */
    uint64_t total = (uint64_t) padk + P;
    uint64_t desired = status + ((1ull << 44) + (total << 22));                /* version_store.h:209-213 */
    uint32_t slot = s_count(status);
    uint64_t *mp = meta_at(leaf, slot);
    uint64_t offset = (uint64_t) n_size(leaf) - s_block(desired);
    uint64_t m = (((uint64_t) ks << 48) | (offset << 32) | cid) | M_CONTROL | M_VISIBLE;  /* record_meta.h:114-120 */
    n_set_status(leaf, desired);
    mp[0] = m;
    mp[2] = 0;                                               /* loc_ptr: host-only indirection, not modelled */
    memcpy(leaf + offset, key, ks);
    memcpy(leaf + offset + padk, payload, P);
    *meta_out = mp;
    return RC_OK;
}

typedef struct { uint64_t *mp; const u8 *key; uint32_t len; } sort_item;
static int sort_cmp(const void *a, const void *b) {
    const sort_item *x = (const sort_item *) a, *y = (const sort_item *) b;
    return so_key_compare(x->key, x->len, y->key, y->len);
}

/* LeafNode::CopyFrom (b_tree.cpp:1486-1545) */
static void leaf_copy_from(so_tree *t, u8 *dst, const u8 *src, sort_item *it, uint32_t n) {
    uint32_t offset = n_size(dst), nrec = 0, P = t->payload_size;
    for (uint32_t i = 0; i < n; i++) {
        uint64_t m = it[i].mp[0];
        if (m == 0 || m_padlen(m) == 0) continue;
        uint32_t total = m_padlen(m) + P;
        offset -= total;
        memcpy(dst + offset, src + m_offset(m), total);
        uint64_t *d = meta_at(dst, nrec);
        d[1] = it[i].mp[1];                                  /* next_ptr carried over */
        d[2] = 0;
        d[0] = m_finalize(offset, m_keylen(m), m_cid(m));
        nrec++;
    }
    uint64_t s = 0;
    s |= ((uint64_t) (n_size(dst) - offset) << 22) & S_BLOCK;
    s |= ((uint64_t) nrec << 44) & S_COUNT;
    n_set_status(dst, s);
    memcpy(dst + OFF_SORTED, &nrec, 4);
}

static u8 *inner_alloc(uint32_t alloc_size) {
    u8 *n = (u8 *) calloc(1, alloc_size);
    memcpy(n + OFF_SIZE, &alloc_size, 4);
    return n;
}
static inline void inner_put(u8 *n, uint32_t idx, uint64_t off, const void *data, uint32_t klen, uint32_t nbytes) {
    *meta_at(n, idx) = m_finalize(off, klen, 0);
    memcpy(n + off, data, nbytes);
}

/* InternalNode(node_size, key, left, right) — single separator (b_tree.cpp:363-389) */
static u8 *inner_new_root(const void *key, uint32_t ks, u8 *left, u8 *right) {
    uint32_t padk = pad8(ks);
    uint32_t alloc = INNER_HDR + padk + 8 + 8 + 2 * META_SZ;                     /* b_tree.h:249-256 */
    u8 *n = inner_alloc(alloc);
    uint32_t two = 2;
    memcpy(n + OFF_SORTED, &two, 4);
    uint64_t off = alloc - 8, l = (uint64_t) (uintptr_t) left, r = (uint64_t) (uintptr_t) right;
    inner_put(n, 0, off, &l, 0, 8);
    off -= padk + 8;
    *meta_at(n, 1) = m_finalize(off, ks, 0);
    memcpy(n + off, key, ks);
    memcpy(n + off + padk, &r, 8);
    return n;
}

/* InternalNode(node_size, src, begin, nr, key, left, right, left_most) (b_tree.cpp:391-497) */
static u8 *inner_new_from(const u8 *src, uint32_t begin, uint32_t nr, const void *key, uint32_t ks, u8 *left,
                          u8 *right, u8 *left_most, uint32_t alloc) {
    u8 *n = inner_alloc(alloc);
    uint32_t padk = pad8(ks), idx = 0;
    uint64_t off = alloc, l = (uint64_t) (uintptr_t) left, r = (uint64_t) (uintptr_t) right;
    int need_new = key != NULL;
    if (left_most) {
        uint64_t lm = (uint64_t) (uintptr_t) left_most;
        off -= 8;
        inner_put(n, 0, off, &lm, 0, 8);
        idx++;
    }
    for (uint32_t i = begin; i < begin + nr; i++) {
        uint64_t m = *meta_at(src, i);
        uint32_t rec = m_padlen(m) + 8;
        const u8 *data = src + m_offset(m);
        if (need_new && so_key_compare(inner_key(src, m), m_keylen(m), key, ks) > 0) {
            uint64_t pm = *meta_at(n, idx - 1);              /* previous entry now points at the left child */
            memcpy(n + m_offset(pm) + m_padlen(pm), &l, 8);
            off -= padk + 8;
            *meta_at(n, idx) = m_finalize(off, ks, 0);
            memcpy(n + off, key, ks);
            memcpy(n + off + padk, &r, 8);
            idx++;
            need_new = 0;
        }
        off -= rec;
        inner_put(n, idx, off, data, m_keylen(m), rec);
        idx++;
    }
    if (need_new) {                                          /* new separator is the right-most entry */
        off -= padk + 8;
        *meta_at(n, idx) = m_finalize(off, ks, 0);
        memcpy(n + off, key, ks);
        memcpy(n + off + padk, &r, 8);
        uint64_t pm = *meta_at(n, idx - 1);
        memcpy(n + m_offset(pm) + m_padlen(pm), &l, 8);
        idx++;
    }
    memcpy(n + OFF_SORTED, &idx, 4);
    return n;
}

/* alloc size of the ranged copy (InternalNode::New, b_tree.h:266-303) */
static uint32_t inner_range_alloc(const u8 *src, uint32_t begin, uint32_t nr, const void *key, uint32_t ks) {
    uint32_t a = INNER_HDR;
    if (begin > 0) a += m_padlen(*meta_at(src, 0)) + 8 + META_SZ;
    for (uint32_t i = begin; i < begin + nr; i++) a += m_padlen(*meta_at(src, i)) + 8 + META_SZ;
    if (key) a += pad8(ks) + 8 + META_SZ;
    return a;
}

/* InternalNode::PrepareForSplit (b_tree.cpp:500-598) */
static u8 *inner_prepare_split(so_tree *t, u8 *node, stack_t_ *st, const void *key, uint32_t ks, u8 *left, u8 *right) {
    uint32_t data_size = n_size(node) + ks + 8 + META_SZ;
    uint32_t new_node_size = INNER_HDR + data_size;
    if (new_node_size < t->split_threshold) {
        uint32_t alloc = n_size(node) + pad8(ks) + 8 + META_SZ;                  /* b_tree.h:230-236 */
        return inner_new_from(node, 0, n_sorted(node), key, ks, left, right, NULL, alloc);
    }
    uint32_t cnt = n_sorted(node), n_left = cnt >> 1;
    uint64_t sm = *meta_at(node, n_left);
    const u8 *sep_key = inner_key(node, sm);
    uint32_t sep_len = m_keylen(sm);
    u8 *sep_child = inner_child(node, sm);
    int cmp = so_key_compare(key, ks, sep_key, sep_len);
    u8 *nl, *nr_;
    if (cmp < 0) {
        nl = inner_new_from(node, 0, n_left, key, ks, left, right, NULL, inner_range_alloc(node, 0, n_left, key, ks));
        nr_ = inner_new_from(node, n_left + 1, cnt - n_left - 1, NULL, 0, NULL, NULL, sep_child,
                             inner_range_alloc(node, n_left + 1, cnt - n_left - 1, NULL, 0));
    } else {
        nl = inner_new_from(node, 0, n_left, NULL, 0, NULL, NULL, NULL, inner_range_alloc(node, 0, n_left, NULL, 0));
        nr_ = inner_new_from(node, n_left + 1, cnt - n_left - 1, key, ks, left, right, sep_child,
                             inner_range_alloc(node, n_left + 1, cnt - n_left - 1, key, ks));
    }
    if (st->n) st->n--;                                      /* stack.Pop() */
    u8 *parent = st->n ? st->f[st->n - 1].node : NULL;
    if (!parent) return inner_new_root(sep_key, sep_len, nl, nr_);
    n_set_status(parent, n_status(parent) | S_FROZEN);
    return inner_prepare_split(t, parent, st, sep_key, sep_len, nl, nr_);
}

/* LeafNode::PrepareForSplit (b_tree.cpp:1558-1690) */
static u8 *leaf_prepare_split(so_tree *t, u8 *leaf, stack_t_ *st) {
    uint32_t count = s_count(n_status(leaf)), P = t->payload_size;
    if (count < 3) return NULL;
    sort_item *v = (sort_item *) malloc(sizeof(sort_item) * count);
    uint32_t n = 0, total = 0;
    for (uint32_t i = 0; i < count; i++) {
        uint64_t *mp = meta_at(leaf, i);
        uint64_t m = mp[0];
        if (m == 0) continue;
        if (m_visible(m) && m_keylen(m) > 0 && m_padlen(m) > 0) {
            v[n].mp = mp; v[n].key = leaf + m_offset(m); v[n].len = m_keylen(m); n++;
            total += m_padlen(m) + P;
        }
    }
    if (total == 0) { free(v); return NULL; }
    qsort(v, n, sizeof(sort_item), sort_cmp);                /* Sorter::Sort, b_tree.h:551-565 */
    int32_t left_size = (int32_t) (total / 2);
    uint32_t nleft = 0;
    for (uint32_t i = 0; i < n; i++) {
        ++nleft;
        left_size -= (int32_t) (m_padlen(v[i].mp[0]) + P);
        if (left_size <= 0) break;
    }
    u8 *l = leaf_new(t, n_size(leaf)), *r = leaf_new(t, n_size(leaf));
    leaf_copy_from(t, l, leaf, v, nleft);
    leaf_copy_from(t, r, leaf, v + nleft, n - nleft);
    /* separator = largest key of the new left leaf: "go left if <=, right if >" (b_tree.cpp:1651-1657) */
    const u8 *sep = v[nleft - 1].key;
    uint32_t sep_len = v[nleft - 1].len;
    u8 *parent = st->n ? st->f[st->n - 1].node : NULL, *res;
    if (!parent) {
        res = inner_new_root(sep, sep_len, l, r);
    } else {
        n_set_status(parent, n_status(parent) | S_FROZEN);
        res = inner_prepare_split(t, parent, st, sep, sep_len, l, r);
    }
    free(v);
    return res;
}

/* BTree::Insert + BTree::FinalizeInsert (b_tree.cpp:1849-2020, 965-977) */
int so_insert(so_tree *t, const void *key, uint32_t ks, const void *payload, uint32_t commit_id) {
    for (;;) {
        stack_t_ st; st.n = 0; st.root = NULL;
        u8 *leaf = traverse_to_leaf(t, &st, key, ks, 1);
        if (!leaf) return RC_NOT_FOUND;
        uint64_t *mp = NULL;
        int rc = leaf_insert(t, leaf, key, ks, payload, commit_id, &mp);
        if (rc == RC_OK) {
            mp[0] = ((mp[0] & ~M_TXN) | commit_id) & ~M_CONTROL;   /* FinalizeForInsert(commit_id), record_meta.h:139-144 */
            return 0;
        }
        if (rc == RC_KEY_EXISTS) return RC_KEY_EXISTS;
        n_set_status(leaf, n_status(leaf) | S_FROZEN);       /* Freeze, b_tree.cpp:124-137 */
        u8 *new_parent = leaf_prepare_split(t, leaf, &st);
        if (!new_parent) return RC_NO_SPACE;                  /* reference would spin; records too large for a leaf */
        u8 *old_parent = st.n ? st.f[--st.n].node : NULL;
        frame_t *top = st.n ? &st.f[--st.n] : NULL;
        if (top) {                                           /* InternalNode::Update: swap the child pointer */
            uint64_t gm = *meta_at(top->node, top->meta_index);
            uint64_t np = (uint64_t) (uintptr_t) new_parent;
            memcpy(top->node + m_offset(gm) + m_padlen(gm), &np, 8);
        } else {
            t->root = new_parent;                            /* ChangeRoot */
        }
        free(leaf);
        t->n_leaves--;
        if (old_parent) free(old_parent);
    }
}

int64_t so_insert_batch(so_tree *t, const void *keys, uint32_t ks, const void *payloads, uint64_t n, uint32_t cid) {
    const u8 *k = (const u8 *) keys, *p = (const u8 *) payloads;
    int64_t ok = 0;
    for (uint64_t i = 0; i < n; i++)
        ok += so_insert(t, k + i * ks, ks, p + i * (uint64_t) t->payload_size, cid) == 0;
    return ok;
}
int64_t so_insert_varlen(so_tree *t, const void *keys, const uint32_t *off, const void *payloads, uint64_t n,
                         uint32_t cid) {
    const u8 *k = (const u8 *) keys, *p = (const u8 *) payloads;
    int64_t ok = 0;
    for (uint64_t i = 0; i < n; i++)
        ok += so_insert(t, k + off[i], off[i + 1] - off[i], p + i * (uint64_t) t->payload_size, cid) == 0;
    return ok;
}

/* LoadYCSBRows, int branch (benchmark/ycsb/ycsb_loader.cpp:144-172) */
int64_t so_load_ycsb_int(so_tree *t, uint64_t begin_rowid, uint64_t end_rowid) {
    u8 *payload = (u8 *) malloc(t->payload_size);
    int64_t ok = 0;
    for (uint64_t rowid = begin_rowid; rowid < end_rowid; rowid++) {
        memset(payload, (int) rowid, t->payload_size);
        uint32_t k = (uint32_t) rowid;
        ok += so_insert(t, &k, 4, payload, 0) == 0;
    }
    free(payload);
    return ok;
}

/* LeafNode::Delete, is_for_update branch: the slot's meta word becomes 0 (b_tree.cpp:1208-1221);
 * BTree::Delete's CheckMerge is a no-op because it is handed an empty stack (b_tree.cpp:2194, 169-173). */
int so_delete(so_tree *t, const void *key, uint32_t ks) {
    u8 *leaf = traverse_to_leaf(t, NULL, key, ks, 1);
    if (!leaf || (n_status(leaf) & S_FROZEN)) return -RC_FROZEN;
    int64_t i = search_record_meta(leaf, key, ks);
    if (i < 0) return -RC_NOT_FOUND;
    meta_at(leaf, (uint32_t) i)[0] = 0;
    return 0;
}

/* TupleHeader bytes (include/vstore/version_store.h:28-46; 56 B): begin_id_@4 comm_id_@8 next_@16 tuple_slot_@40 */
enum { TH_SIZE = 56, TH_BEGIN = 4, TH_COMM = 8, TH_NEXT = 16, TH_SLOT = 40 };

/* committed update, final state (src/concurrency/transaction_manager.cpp:610-676, ephemeral_pool.cpp:17-44) */
int so_commit_update(so_tree *t, const void *key, uint32_t ks, const void *new_payload, uint32_t commit_id) {
    u8 *leaf = traverse_to_leaf(t, NULL, key, ks, 1);
    if (!leaf) return -1;
    int64_t i = search_record_meta(leaf, key, ks);
    if (i < 0) return -2;
    uint64_t *mp = meta_at(leaf, (uint32_t) i);
    uint32_t klen = m_keylen(mp[0]), P = t->payload_size;
    u8 *rec = leaf + m_offset(mp[0]);
    u8 *mem = (u8 *) calloc(1, TH_SIZE + klen + P);
    if (t->n_chain == t->cap_chain) {
        t->cap_chain = t->cap_chain ? t->cap_chain * 2 : 1024;
        t->chain_allocs = (void **) realloc(t->chain_allocs, t->cap_chain * sizeof(void *));
    }
    t->chain_allocs[t->n_chain++] = mem;
    u8 *slot = mem + TH_SIZE;
    memcpy(slot, rec, klen);
    memcpy(slot + klen, rec + m_padlen(mp[0]), P);
    uint32_t begin = m_cid(mp[0]);
    uint64_t slot_addr = (uint64_t) (uintptr_t) slot, next = mp[1];
    memcpy(mem + TH_BEGIN, &begin, 4);
    memcpy(mem + TH_COMM, &commit_id, 4);
    memcpy(mem + TH_NEXT, &next, 8);
    memcpy(mem + TH_SLOT, &slot_addr, 8);
    memcpy(rec + m_padlen(mp[0]), new_payload, P);
    mp[0] = ((mp[0] & ~M_TXN) | commit_id) & ~M_CONTROL;     /* FinalizeForUpdate, record_meta.h:146-151 */
    mp[1] = (uint64_t) (uintptr_t) mem;
    return 0;
}

/* n committed single-row updates in a row: update i rewrites column 1 (payload bytes [100,200) clipped to
 * the payload) with (uint8)(first_cid + i) and commits with id first_cid + i — the shape of the YCSB
 * update transaction (benchmark/ycsb/ycsb_mixed.cpp:38-80) after commit. Returns updates applied. */
int64_t so_commit_update_batch(so_tree *t, const void *keys, uint32_t ks, uint64_t n, uint32_t first_cid) {
    const u8 *k = (const u8 *) keys;
    uint32_t P = t->payload_size;
    u8 *np_ = (u8 *) malloc(P);
    int64_t ok = 0;
    for (uint64_t i = 0; i < n; i++) {
        u8 *leaf = traverse_to_leaf(t, NULL, k + i * ks, ks, 1);
        int64_t s = leaf ? search_record_meta(leaf, k + i * ks, ks) : -1;
        if (s < 0) continue;
        uint64_t m = meta_at(leaf, (uint32_t) s)[0];
        memcpy(np_, leaf + m_offset(m) + m_padlen(m), P);
        uint32_t lo = P > 100 ? 100 : 0, hi = P > 200 ? 200 : P;
        memset(np_ + lo, (int) ((first_cid + i) & 0xFF), hi - lo);
        ok += so_commit_update(t, k + i * ks, ks, np_, first_cid + (uint32_t) i) == 0;
    }
    free(np_);
    return ok;
}

/* ------------------------------------------------------------------------------------------ */
/* a10: visibility rule + version chain (include/execute/executor.h:383-446, 465-534)          */
/* UB branches at HEAD are given defined behaviour: emit the hit once.       */
/* ------------------------------------------------------------------------------------------ */
static int visible_version(const u8 *leaf, const uint64_t *mp, uint32_t ks, uint32_t P, uint32_t read_ts,
                           uint32_t *ts_out, u8 *key_out, u8 *payload_out) {
    uint64_t m = mp[0];
    uint32_t cid = m_cid(m);
    const u8 *rec = leaf + m_offset(m);
    if (read_ts >= cid) {
        if (ts_out) *ts_out = cid;
        if (key_out) memcpy(key_out, rec, ks);
        if (payload_out) memcpy(payload_out, rec + m_padlen(m), P);
        return 1;
    }
    uint64_t next = mp[1];
    while (next != 0 && next != UINT64_MAX) {
        const u8 *hd = (const u8 *) (uintptr_t) next;
        uint32_t begn, comm;
        memcpy(&begn, hd + TH_BEGIN, 4);
        memcpy(&comm, hd + TH_COMM, 4);
        if (begn == 0 || comm == 0) return 0;                /* INVALID_CID stops the walk */
        if (read_ts >= begn && read_ts <= comm) {
            uint64_t sa; memcpy(&sa, hd + TH_SLOT, 8);
            const u8 *slot = (const u8 *) (uintptr_t) sa;
            if (ts_out) *ts_out = begn;
            if (key_out) memcpy(key_out, slot, ks);
            if (payload_out) memcpy(payload_out, slot + ks, P);
            return 1;
        }
        memcpy(&next, hd + TH_NEXT, 8);
    }
    return 0;
}

/* a5: BTree::Read (b_tree.cpp:2066-2129), quiesced tree (no IsInserting records) */
static inline int probe_one(so_tree *t, const u8 *key, uint32_t ks, uint32_t read_ts, uint32_t *ts, u8 *key_out,
                            u8 *payload_out) {
    u8 *leaf = traverse_to_leaf(t, NULL, key, ks, 1);
    if (!leaf) return 0;
    int64_t i = search_record_meta(leaf, key, ks);           /* LeafNode::Read, b_tree.cpp:1042-1051 */
    if (i < 0) return 0;
    return visible_version(leaf, meta_at(leaf, (uint32_t) i), ks, t->payload_size, read_ts, ts, key_out, payload_out);
}

int so_probe(so_tree *t, const void *keys, uint32_t ks, uint64_t n, uint32_t read_ts, uint8_t *found, uint32_t *ts,
             void *key_out, void *payload_out) {
    const u8 *k = (const u8 *) keys;
    uint32_t P = t->payload_size;
    for (uint64_t i = 0; i < n; i++) {
        int f = probe_one(t, k + i * ks, ks, read_ts, ts ? ts + i : NULL, key_out ? (u8 *) key_out + i * ks : NULL,
                          payload_out ? (u8 *) payload_out + i * (uint64_t) P : NULL);
        if (found) found[i] = (uint8_t) f;
        if (!f) {
            if (ts) ts[i] = 0;
            if (key_out) memset((u8 *) key_out + i * ks, 0, ks);
            if (payload_out) memset((u8 *) payload_out + i * (uint64_t) P, 0, P);
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* a9: LeafNode::RangeScanBySize (b_tree.cpp:1261-1315) and a8: Iterator (b_tree.h:883-941)     */
/* ------------------------------------------------------------------------------------------ */
typedef struct { const u8 *leaf; const uint64_t *mp; const u8 *key; uint32_t len; } scan_item;
typedef struct { scan_item *v; uint32_t n, head, cap; } scan_buf;

static int scan_item_cmp(const void *a, const void *b) {
    const scan_item *x = (const scan_item *) a, *y = (const scan_item *) b;
    return so_key_compare(x->key, x->len, y->key, y->len);
}

static void leaf_range_scan(const u8 *leaf, const void *key1, uint32_t size1, uint32_t to_scan, scan_buf *out) {
    out->n = 0; out->head = 0;
    if (to_scan == 0) return;
    uint32_t count = s_count(n_status(leaf));
    for (uint32_t i = 0; i < count; i++) {
        if (out->n > to_scan) break;                         /* slot-order early break, BEFORE the sort (Q5) */
        const uint64_t *mp = meta_at(leaf, i);
        uint64_t m = mp[0];
        if (!m_visible(m)) continue;
        const u8 *k = leaf + m_offset(m);
        if (so_key_compare(key1, size1, k, m_keylen(m)) <= 0) {
            if (out->n == out->cap) {
                out->cap = out->cap ? out->cap * 2 : 128;
                out->v = (scan_item *) realloc(out->v, out->cap * sizeof(scan_item));
            }
            scan_item it = { leaf, mp, k, m_keylen(m) };
            out->v[out->n++] = it;
        }
    }
    qsort(out->v, out->n, sizeof(scan_item), scan_item_cmp);
}

/* executor loop over Iterator::GetNext (executor.h:459-535 with visibility, :626-638 without) */
static uint32_t scan_one(so_tree *t, scan_buf *buf, const u8 *start, uint32_t ks, uint32_t scan_size, uint32_t read_ts,
                         int apply_visibility, uint32_t *ts_out, u8 *key_out, u8 *payload_out, uint32_t kstride,
                         uint32_t *lens_out) {
    uint32_t P = t->payload_size, remaining = scan_size, cnt = 0;
    u8 *leaf = traverse_to_leaf(t, NULL, start, ks, 1);      /* Iterator ctor, b_tree.h:885-891 */
    leaf_range_scan(leaf, start, ks, scan_size, buf);
    for (uint32_t scanned = 0; scanned < scan_size; scanned++) {
        if (buf->head == buf->n || remaining == 0) continue; /* GetNext returns nullptr */
        remaining -= 1;
        scan_item cur = buf->v[buf->head++];
        if (buf->n - (buf->head - 1) == 1) {                 /* it was the only buffered record: hop */
            u8 *nl = traverse_to_leaf(t, NULL, cur.key, cur.len, 0);
            buf->n = buf->head = 0;
            if (nl) {
                leaf_range_scan(nl, cur.key, cur.len, remaining, buf);
                if (buf->n && so_key_compare(buf->v[0].key, buf->v[0].len, cur.key, cur.len) == 0)
                    buf->n = buf->head = 0;                  /* "hit the same record": scan ends */
            }
        }
        /* IndexScanExecutor range branch, executor.h:465-534: the row is returned iff read_id >= the commit id of
         * its in-leaf version. The else-branch reads RecordMeta::next_tuple_ptr, which is 0 for every record the
         * iterator hands out (RecordMeta ctor record_meta.h:240; LeafNode::RangeScanBySize b_tree.cpp:1290-1297
         * never sets it; only BTree::Read does, b_tree.cpp:2099/2116): no chain walk ever happens in a range scan,
         * a too-new row is dropped. Verified on the unmodified executor (tests/cpp/dropin_test.cpp). */
        int f = 1;
        if (apply_visibility && read_ts < m_cid(cur.mp[0])) {
            f = 0;
        } else {
            uint64_t m = cur.mp[0];
            if (ts_out) ts_out[cnt] = m_cid(m);
            if (lens_out) {                                  /* variable-length keys: key bytes + their length */
                lens_out[cnt] = cur.len;
                if (key_out) memcpy(key_out + (uint64_t) cnt * kstride, cur.key, cur.len < kstride ? cur.len : kstride);
            } else if (key_out) {
                memcpy(key_out + (uint64_t) cnt * ks, cur.key, ks);
            }
            if (payload_out) memcpy(payload_out + (uint64_t) cnt * P, cur.leaf + m_offset(m) + m_padlen(m), P);
        }
        cnt += (uint32_t) f;
    }
    return cnt;
}

int so_scan(so_tree *t, const void *start_keys, uint32_t ks, const uint32_t *scan_sizes, uint64_t n,
            uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts, void *key_out,
            uint32_t *ts_out, void *payload_out) {
    const u8 *k = (const u8 *) start_keys;
    uint32_t P = t->payload_size;
    scan_buf buf = { 0 };
    for (uint64_t i = 0; i < n; i++) {
        uint64_t base = i * (uint64_t) out_stride;
        row_counts[i] = scan_one(t, &buf, k + i * ks, ks, scan_sizes[i], read_ts, apply_visibility,
                                 ts_out ? ts_out + base : NULL, key_out ? (u8 *) key_out + base * ks : NULL,
                                 payload_out ? (u8 *) payload_out + base * P : NULL, ks, NULL);
    }
    free(buf.v);
    return 0;
}

/* the same loop for tables with keys of different lengths (test/testing_btree.cpp's std::to_string keys): start key i
 * = keys[off[i] .. off[i+1]); returned keys go to key_out in slots of key_stride bytes, their lengths to key_lens */
int so_scan_varlen(so_tree *t, const void *keys, const uint32_t *off, const uint32_t *scan_sizes, uint64_t n,
                   uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts, void *key_out,
                   uint32_t key_stride, uint32_t *key_lens, uint32_t *ts_out, void *payload_out) {
    const u8 *k = (const u8 *) keys;
    uint32_t P = t->payload_size;
    scan_buf buf = { 0 };
    for (uint64_t i = 0; i < n; i++) {
        uint64_t base = i * (uint64_t) out_stride;
        row_counts[i] = scan_one(t, &buf, k + off[i], off[i + 1] - off[i], scan_sizes[i], read_ts, apply_visibility,
                                 ts_out ? ts_out + base : NULL, key_out ? (u8 *) key_out + base * key_stride : NULL,
                                 payload_out ? (u8 *) payload_out + base * P : NULL, key_stride, key_lens + base);
    }
    free(buf.v);
    return 0;
}

/* BTree::Read for keys of different lengths: key i = keys[off[i] .. off[i+1]) */
int so_probe_varlen(so_tree *t, const void *keys, const uint32_t *off, uint64_t n, uint32_t read_ts, uint8_t *found,
                    uint32_t *ts, void *payload_out) {
    const u8 *k = (const u8 *) keys;
    uint32_t P = t->payload_size;
    for (uint64_t i = 0; i < n; i++) {
        int f = probe_one(t, k + off[i], off[i + 1] - off[i], read_ts, ts ? ts + i : NULL, NULL,
                          payload_out ? (u8 *) payload_out + i * P : NULL);
        found[i] = (uint8_t) f;
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* in-order walk: leaves + separators (executor.h:592-609 order)                               */
/* ------------------------------------------------------------------------------------------ */
typedef struct { const void **leaves, **sep_keys; uint32_t *sep_lens; uint64_t n, cap; } walk_ctx;
static void walk_rec(const u8 *n, walk_ctx *c) {
    if (n_is_leaf(n)) {
        if (c->n < c->cap && c->leaves) c->leaves[c->n] = n;
        c->n++;
        return;
    }
    uint32_t cnt = n_sorted(n);
    for (uint32_t i = 0; i < cnt; i++) {
        uint64_t m = *meta_at(n, i);
        if (i > 0) {                                         /* separator i sits between child i-1 and child i */
            uint64_t at = c->n - 1;
            if (at < c->cap && c->sep_keys) { c->sep_keys[at] = n + m_offset(m); c->sep_lens[at] = m_keylen(m); }
        }
        walk_rec(inner_child(n, m), c);
    }
}
uint64_t so_walk(const void *root, const void **leaves, const void **sep_keys, uint32_t *sep_lens, uint64_t cap) {
    walk_ctx c = { leaves, sep_keys, sep_lens, 0, cap };
    walk_rec((const u8 *) root, &c);
    return c.n;
}

/* ------------------------------------------------------------------------------------------ */
/* timed legs                                                                                  */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    so_tree *t; const u8 *keys; uint32_t ks, read_ts, scan_size; uint64_t b, e, out; int vis, pin, core, is_scan;
    pthread_barrier_t *bar;
} job_t;

static void *job_main(void *arg) {
    job_t *j = (job_t *) arg;
    if (j->pin) { cpu_set_t s; CPU_ZERO(&s); CPU_SET(j->core, &s); pthread_setaffinity_np(pthread_self(), sizeof(s), &s); }
    uint32_t P = j->t->payload_size;
    uint64_t acc = 0;
    if (!j->is_scan) {
        u8 *pay = (u8 *) malloc(P), *key = (u8 *) malloc(j->ks);
        uint32_t ts;
        pthread_barrier_wait(j->bar);
        for (uint64_t i = j->b; i < j->e; i++) acc += probe_one(j->t, j->keys + i * j->ks, j->ks, j->read_ts, &ts, key, pay);
        free(pay); free(key);
    } else {
        u8 *pay = (u8 *) malloc((size_t) P * j->scan_size), *key = (u8 *) malloc((size_t) j->ks * j->scan_size);
        uint32_t *ts = (uint32_t *) malloc(4 * (size_t) j->scan_size);
        scan_buf buf = { 0 };
        pthread_barrier_wait(j->bar);
        for (uint64_t i = j->b; i < j->e; i++)
            acc += scan_one(j->t, &buf, j->keys + i * j->ks, j->ks, j->scan_size, j->read_ts, j->vis, ts, key, pay, j->ks, NULL);
        free(pay); free(key); free(ts); free(buf.v);
    }
    j->out = acc;
    pthread_barrier_wait(j->bar);
    return NULL;
}

static double run_jobs(so_tree *t, const void *keys, uint32_t ks, uint64_t n, uint32_t read_ts, uint32_t scan_size,
                       int vis, int threads, int pin, int is_scan, uint64_t *out) {
    pthread_t *th = (pthread_t *) malloc(sizeof(pthread_t) * threads);
    job_t *jobs = (job_t *) calloc(threads, sizeof(job_t));
    pthread_barrier_t bar;
    pthread_barrier_init(&bar, NULL, threads + 1);
    for (int w = 0; w < threads; w++) {
        job_t j = { t, (const u8 *) keys, ks, read_ts, scan_size, n * w / threads, n * (w + 1) / threads, 0, vis, pin, w,
                    is_scan, &bar };
        jobs[w] = j;
        pthread_create(&th[w], NULL, job_main, &jobs[w]);
    }
    struct timespec t0, t1;
    pthread_barrier_wait(&bar);
    clock_gettime(CLOCK_MONOTONIC, &t0);
    pthread_barrier_wait(&bar);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    uint64_t tot = 0;
    for (int w = 0; w < threads; w++) { pthread_join(th[w], NULL); tot += jobs[w].out; }
    if (out) *out = tot;
    pthread_barrier_destroy(&bar);
    free(th); free(jobs);
    return (double) (t1.tv_sec - t0.tv_sec) + 1e-9 * (double) (t1.tv_nsec - t0.tv_nsec);
}

double so_time_probe(so_tree *t, const void *keys, uint32_t ks, uint64_t n, uint32_t read_ts, int threads, int pin,
                     uint64_t *hits_out) {
    return run_jobs(t, keys, ks, n, read_ts, 0, 0, threads, pin, 0, hits_out);
}
double so_time_scan(so_tree *t, const void *start_keys, uint32_t ks, uint32_t scan_size, uint64_t n, uint32_t read_ts,
                    int apply_visibility, int threads, int pin, uint64_t *rows_out) {
    return run_jobs(t, start_keys, ks, n, read_ts, scan_size, apply_visibility, threads, pin, 1, rows_out);
}

"""Host-side (numpy) reader of one Stage BzTree leaf image, mirroring csrc/node_layout.h — used by tests and tools to
state expectations about leaf contents (slot order, visibility) without going through any search code."""
import numpy as np

OFF_STATUS, OFF_META, META_SIZE = 32, 40, 24


def parse_leaf(img, key_size, payload_size):
    """(keys [n,K], payloads [n,P]) of the VISIBLE records of a leaf image, in SLOT order
    (RecordMetadata layout: include/vstore/record_meta.h:30-70; status word: version_store.h:176-217)."""
    img = np.ascontiguousarray(img, np.uint8)
    status = int(img[OFF_STATUS:OFF_STATUS + 8].view(np.uint64)[0])
    count = (status >> 44) & 0xFFFF
    meta = img[OFF_META:OFF_META + count * META_SIZE].reshape(count, META_SIZE)[:, :8].copy().view(np.uint64).ravel()
    vis = (meta >> np.uint64(62)) & np.uint64(1)
    off = ((meta >> np.uint64(32)) & np.uint64(0xFFFF)).astype(np.int64)[vis == 1]
    pad = (key_size + 7) // 8 * 8
    keys = np.stack([img[o:o + key_size] for o in off]) if len(off) else np.zeros((0, key_size), np.uint8)
    pays = np.stack([img[o + pad:o + pad + payload_size] for o in off]) if len(off) else np.zeros((0, payload_size), np.uint8)
    return keys, pays

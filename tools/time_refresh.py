"""Times stage_gpu_snapshot_refresh_from_tree against a full stage_gpu_snapshot_from_tree on the bench table
(10 M rows, 1 M committed updates) after further host-side changes. Test infrastructure (uses oracle/)."""
import json, sys, time
sys.path.insert(0, '.')
import numpy as np
import stage_indexorganized_b200 as sg
import bench

sg.init(0)
IN_PLACE = "--in-place" in sys.argv          # stage_gpu_snapshot_refresh_in_place (rows shifted inside the payload buffer)
args = [a for a in sys.argv[1:] if not a.startswith("--")]
ROWS = int(args[0]) if args else 10_000_000
t, kind = bench.build_host_table(ROWS, ROWS // 10)
if IN_PLACE:
    sg.set_row_slack(0.05)
t0 = time.time(); snap = sg.Snapshot.from_tree(t.root, 4, 1000); full_s = time.time() - t0
out = {"rows": ROWS, "table": kind, "mode": "in place" if IN_PLACE else "two tables", "from_tree_s": full_s,
       "leaves": snap.info().n_leaves, "device_GB": snap.info().device_bytes / 1e9, "cases": []}
cid = 10 + ROWS // 10 + 100
rng = np.random.default_rng(5)
for what, n_upd, n_ins in (("no change", 0, 0), ("1 k updates", 1000, 0), ("10 k updates", 10_000, 0),
                           ("100 k updates", 100_000, 0), ("10 k inserts in one key region", 0, 10_000),
                           ("10 k updates + 10 k inserts", 10_000, 10_000)):
    if n_upd:
        k = rng.integers(0, ROWS, n_upd).astype(np.uint32).view(np.uint8).reshape(-1, 4)
        t.commit_update_batch(k, cid); cid += n_upd
    if n_ins:
        base = 0x30000009 + 0x01000000 * len(out["cases"])
        ik = (base + 256 * np.arange(n_ins)).astype(np.uint32)
        pay = np.repeat((ik >> 8).astype(np.uint8)[:, None], 1000, axis=1)
        assert t.insert(ik.view(np.uint8).reshape(-1, 4), pay, cid) == n_ins; cid += 1
    if IN_PLACE:
        t0 = time.time(); st = snap.refresh_in_place(t.root); dt = time.time() - t0; nxt = snap
    else:
        t0 = time.time(); nxt, st = snap.refresh_from_tree(t.root); dt = time.time() - t0
    out["cases"].append({"change": what, "refresh_s": dt, "leaves_total": st.leaves_total,
                         "leaves_transformed": st.leaves_transformed, "host_MB_staged": st.host_bytes_staged / 1e6,
                         "speedup_vs_from_tree": full_s / dt})
    print(out["cases"][-1], flush=True)
    if nxt is not snap:
        snap.close(); snap = nxt
# the refreshed snapshot answers like the tree
pk = rng.integers(0, ROWS + 10, 50_000).astype(np.uint32).view(np.uint8).reshape(-1, 4)
for ts in (0, 10 + ROWS // 20, cid + 5):
    g, w = snap.probe(pk, ts), t.probe(pk, ts)
    assert all(np.array_equal(a, b) for a, b in zip(g, w) if a is not None and b is not None), ts
out["parity_after_last_refresh"] = True
t0 = time.time(); kept, dropped = snap.compact_versions(); out["compact_versions"] = {"seconds": time.time() - t0, "kept": kept, "dropped": dropped}
out["device_GB_end"] = snap.info().device_bytes / 1e9
print(json.dumps(out))

"""Prototype (CPU, numpy) of the event replay that turns the clean lattice numbering into the
reference's numbering with hash-growth orphans; checked against Oracle(faithful=True).
Development aid for csrc/pl_grow.cu -- not product code."""
import sys, bisect
import numpy as np
sys.path.insert(0, ".")
from oracle.lattice_oracle import Oracle

MASK64 = (1 << 64) - 1

def ref_hash(key):
    k = 0
    for x in key:
        k = (k + (int(x) & MASK64)) & MASK64
        k = (k * 2531011) & MASK64
    return k

def replay(vid_clean, keys_clean, d):
    E = vid_clean.size
    M = keys_clean.shape[0]
    # first touch entry of each clean vertex (monotone)
    first_e = np.full(M, E, np.int64)
    np.minimum.at(first_e, vid_clean, np.arange(E))
    assert np.all(np.diff(first_e) > 0)
    extras = []            # sorted creation entries of extra copies
    special = {}           # v -> dict(live=copy_entry|None, parked=copy_entry|None, segs=[(start_entry, copy_entry)])
    def filled_after(e):   # number of vertices created by entries <= e
        return int(np.searchsorted(first_e, e, side="right")) + bisect.bisect_right(extras, e)
    def first_reach(T):    # min e with filled_after(e) >= T, or None
        lo, hi = -1, E - 1
        if filled_after(hi) < T: return None
        while hi - lo > 1:
            mid = (lo + hi) // 2
            if filled_after(mid) >= T: hi = mid
            else: lo = mid
        return hi
    def next_touch(v, after):
        idx = np.nonzero(vid_clean[after + 1:] == v)[0]
        return None if idx.size == 0 else after + 1 + int(idx[0])
    C = 1 << 15
    pending_second = None   # (v, candidate entry) not yet confirmed
    blur_trigger = None
    t_prev = -1
    while True:
        T = C // 2 - 1
        ec = first_reach(T)
        if ec is None:
            break
        t = ec + 1
        # resolve a pending second-touch candidate against this trigger
        if pending_second is not None:
            v2, e2 = pending_second
            if e2 >= t:
                extras.remove(e2)
                ec = first_reach(T); t = ec + 1      # unchanged by construction, recompute for safety
                special[v2]["segs"] = [s for s in special[v2]["segs"] if s[0] != e2]
                special[v2]["live"] = None
            pending_second = None
        if t >= E:
            blur_trigger = C
            # grow happens at the first blur lookup: parked copies flip
            for s in special.values():
                if s["parked"] is not None:
                    s["live"] = s["parked"]; s["parked"] = None
            C *= 2
            break
        # grow at lookup t
        for s in special.values():
            if s["parked"] is not None:
                s["live"] = s["parked"]; s["segs"].append((t, s["parked"])); s["parked"] = None
        Cold, C = C, C * 2
        v = int(vid_clean[t])
        key = keys_clean[v]
        H = ref_hash(key)
        if (H % Cold) != (H % C):
            is_new = first_e[v] == t and v not in special
            s = special.setdefault(v, dict(live=(None if is_new else int(first_e[v])), parked=None,
                                           segs=([] if is_new else [(int(first_e[v]), int(first_e[v]))])))
            if v in special and not is_new and s["live"] is None:
                pass
            if not is_new:
                extras.append(t); extras.sort()
            s["parked"] = t
            s["segs"].append((t, t))                       # entry t itself uses the parked copy
            if s["live"] is not None:
                s["segs"].append((t + 1, s["live"]))       # later lookups find the old live copy again
            else:
                e2 = next_touch(v, t)
                if e2 is not None:
                    extras.append(e2); extras.sort()
                    s["segs"].append((e2, e2)); s["live"] = e2
                    pending_second = (v, e2)
        else:
            # normal lookup; if the key is special and currently unfindable, it creates a true copy
            if v in special and special[v]["live"] is None:
                raise NotImplementedError("trigger on an unfindable special key")
    return first_e, extras, special, blur_trigger, C

def apply(vid_clean, keys_clean, d, res):
    first_e, extras, special, blur_trigger, Cfinal = res
    E = vid_clean.size
    M = keys_clean.shape[0]
    ex = np.array(sorted(extras), np.int64)
    shift_at = lambda e: np.searchsorted(ex, e, side="left")   # extras created strictly before e
    new_id_clean = np.arange(M) + shift_at(first_e)
    def copy_id(entry):   # id of the copy created at `entry`
        return int(np.searchsorted(first_e, entry, side="left") + np.searchsorted(ex, entry, side="left"))
    vid = new_id_clean[vid_clean].astype(np.int32)
    Mf = M + ex.size
    keys = np.zeros((Mf, d), np.int16)
    keys[new_id_clean] = keys_clean
    live_final = {}
    for v, s in special.items():
        segs = sorted(s["segs"])
        ent = np.nonzero(vid_clean == v)[0]
        starts = [x[0] for x in segs]
        for e in ent:
            k = bisect.bisect_right(starts, e) - 1
            vid[e] = copy_id(segs[k][1])
        for _, c in segs:
            keys[copy_id(c)] = keys_clean[v]
        live_final[v] = copy_id(s["live"]) if s["live"] is not None else None
    return vid, keys, Mf, live_final

if __name__ == "__main__":
    rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
    clean, faith = Oracle(False), Oracle(True)
    for trial in range(int(sys.argv[2]) if len(sys.argv) > 2 else 10):
        d = int(rng.choice([2, 3, 5]))
        n = int(rng.integers(5000, 120000))
        sc = float(rng.choice([8.0, 16.0, 64.0])) * (1 if d == 5 else 12)
        ref = (rng.random((n, d)) * sc).astype(np.float32)
        a = clean.state(ref); b = faith.state(ref)
        res = replay(a.vid.reshape(-1), a.keys, d)
        vid, keys, Mf, live = apply(a.vid.reshape(-1), a.keys, d, res)
        ok = (Mf == b.M) and np.array_equal(vid, b.vid.reshape(-1)) and np.array_equal(keys, b.keys)
        print(f"d={d} n={n} sc={sc} Mclean={a.M} Mref={b.M} extras={len(res[1])} special={len(res[2])} "
              f"blur_trigger={res[3]} ok={ok}")
        if not ok:
            bad = np.nonzero(vid != b.vid.reshape(-1))[0]
            print("  first bad entries", bad[:10], vid[bad[:5]], b.vid.reshape(-1)[bad[:5]])


def neighbours(keys, Mf, d, key_to_live, blur_trigger):
    nbr = np.full((d + 1, Mf, 2), -1, np.int32)
    for j in range(d + 1):
        for v in range(Mf):
            k = keys[v].astype(np.int64)
            n1 = k + 1; n2 = k - 1
            if j < d:
                n1[j] = k[j] - d; n2[j] = k[j] + d
            for s, nk in ((0, n1), (1, n2)):
                nk16 = nk.astype(np.int16)
                got = key_to_live.get(nk16.tobytes(), -1)
                if blur_trigger is not None and j == 0 and v == 0 and s == 0:
                    H = ref_hash(nk16)
                    if (H % blur_trigger) != (H % (2 * blur_trigger)):
                        got = -1
                nbr[j, v, s] = got
    return nbr


def full_check(ref, d, verbose=True):
    clean, faith = Oracle(False), Oracle(True)
    a = clean.state(ref); b = faith.state(ref)
    res = replay(a.vid.reshape(-1), a.keys, d)
    vid, keys, Mf, live = apply(a.vid.reshape(-1), a.keys, d, res)
    ok = (Mf == b.M) and np.array_equal(vid, b.vid.reshape(-1)) and np.array_equal(keys, b.keys)
    nbr_ok = None
    if ok and Mf < 60000:
        first_e, extras, special, blur_trigger, _ = res
        ex = np.array(sorted(extras), np.int64)
        new_id_clean = np.arange(a.M) + np.searchsorted(ex, first_e, side="left")
        k2l = {a.keys[v].tobytes(): int(new_id_clean[v]) for v in range(a.M)}
        for v, lv in live.items():
            if lv is None: k2l.pop(a.keys[v].tobytes())
            else: k2l[a.keys[v].tobytes()] = lv
        nbr = neighbours(keys, Mf, d, k2l, blur_trigger)
        nbr_ok = np.array_equal(nbr, b.nbr)
    if verbose:
        print(f"d={d} n={ref.shape[0]} Mclean={a.M} Mref={b.M} extras={len(res[1])} special={len(res[2])} "
              f"blur_trigger={res[3]} ok={ok} nbr_ok={nbr_ok}")
    return ok and nbr_ok is not False

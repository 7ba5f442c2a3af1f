"""TEST INFRASTRUCTURE ONLY — CPU restatement of VOXEL_SLAM::thd_globalmapping (voxelslam.cpp:3018-3141) for stored sessions,
chained over the oracle's HBA_add_edge (oracle/capi.cpp: orc_hba_add_edge).  Pure-Python control flow, small cases.

schedule_literal() walks the reference's while-loop state machine variable for variable (smp_mp, buf_base, localID, smp_flag,
total_ba) with every keyframe already queued and gba_flag raised once the last session has been consumed — the second,
independent coding vx_hba_schedule is checked against."""
import numpy as np

import oracle_bind as ob


def schedule_literal(session_sizes, wdsize=10, mgsize=5):
    """-> list of (session, [keyframe indices inside the session]) in the order HBA_add_edge is called at :3096."""
    windows = []
    n_sess = len(session_sizes)
    smp_mp, buf_base = 0, 0
    localID = []
    gba_size = sum(session_sizes)          # gba_flag = 1 with gba_size = all keyframes of the last session consumed
    consumed = 0
    while True:
        smp_flag = 1 if smp_mp + 1 < n_sess else 0                     # :3046-3048
        size = session_sizes[smp_mp]
        total_ba = 0
        if smp_mp >= n_sess - 1 and consumed >= gba_size:              # :3054 (gba_flag, cnct_map.back(), gba_size <= buf_base)
            total_ba = 1
        elif size <= buf_base:                                          # :3059
            if smp_flag == 0:
                raise AssertionError("unreachable: the last session ends through total_ba")
        else:                                                           # :3074-3083
            localID.append(buf_base)
            buf_base += 1; consumed += 1
            if len(localID) < wdsize:
                continue
        if localID:
            windows.append((smp_mp, list(localID)))                    # :3086-3099
        if total_ba == 1:                                               # :3102-3125
            return windows
        elif smp_flag == 1 and size <= buf_base:                        # :3127-3133
            smp_mp += 1; buf_base = 0; localID = []
        else:                                                           # :3135-3138
            del localID[:mgsize]


def hba_global(clouds, kf_session, kf_x0, local, gba, max_layer, total_max_iter=1, kf_x=None, kf_id=None, cnct_map=None, wdsize=10, mgsize=5):
    """clouds: one (n,3) float32 array per keyframe; kf_x0 / kf_x: (R (n,3,3), p (n,3)).  Mirrors pkg.hba_global's result."""
    kf_session = np.asarray(kf_session)
    sizes = np.bincount(kf_session).tolist()
    first = np.concatenate([[0], np.cumsum(sizes)])
    R0, p0 = kf_x0
    Rx, px = kf_x if kf_x is not None else kf_x0
    wins = schedule_literal(sizes, wdsize, mgsize)
    edges1, sub_clouds, sub_kf = [], [], []
    for (s, ids) in wins:
        kfs = [first[s] + i for i in ids]
        kid = [int(kf_id[k]) if kf_id is not None else int(k - first[s]) for k in kfs]
        r = ob.hba_add_edge([clouds[k] for k in kfs], (R0[kfs], p0[kfs]), local, gba, max_layer, 1, thread_num=2, ids=kid, mps=[s] * len(kfs), want_cloud=True)
        assert r["n_edges"] >= 0, "Too Less Voxel in a bottom window"
        edges1.append(r["edges"])
        m = r["merged"]
        sub_clouds.append(np.stack([m["x"], m["y"], m["z"]], axis=1).astype(np.float32))
        sub_kf.append(kfs[0])
    cn = set(range(len(sizes))) if cnct_map is None else set(int(c) for c in cnct_map)
    top = [w for w, (s, _) in enumerate(wins) if s in cn]
    out = dict(edges1=np.concatenate(edges1) if edges1 else np.zeros(0, dtype=ob.edge_dtype), windows=wins, sub_clouds=sub_clouds,
               top_kf=np.array([sub_kf[w] for w in top], dtype=np.int32))
    tk = out["top_kf"]
    if len(top) >= 2 and total_max_iter > 0:
        tid = [int(kf_id[k]) if kf_id is not None else int(k - first[kf_session[k]]) for k in tk]
        r = ob.hba_add_edge([sub_clouds[w] for w in top], (Rx[tk], px[tk]), local, gba, max_layer, total_max_iter, thread_num=5, ids=tid,
                            mps=[int(kf_session[k]) for k in tk])
        assert r["n_edges"] >= 0, "Too Less Voxel at the top level"
        out["edges2"] = r["edges"]; out["top_poses"] = r["poses"]; out["top"] = r
    else:
        out["edges2"] = np.zeros(0, dtype=ob.edge_dtype); out["top_poses"] = ob.as_poses((Rx[tk], px[tk]))
    return out

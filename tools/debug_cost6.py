import sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import make_golden_inputs as gi
from oracle_lib import Oracle, Reference
orc, ref = Oracle(), Reference()
W, H = gi.W, gi.H
for dof6 in (False, True):
    case, obs, poses, pmm, lab, tot, stride = gi._cost_inputs(orc, dof6)
    s = case.scene
    for src_name, src in (("scene", s.source_depth), ("zeros", np.zeros_like(s.source_depth))):
        q = ref.render_cost(s.tris, s.tris_model_count, poses, pmm, W, H, case.proj, src, s.source_color,
                            s.source_label if dof6 else None, lab, stride, 100, case.cx, case.cy, case.fx, case.fy, obs["aos"],
                            obs["rgb"], obs["label"] if dof6 else None, tot, 2 if dof6 else 0, True, 0.01 if dof6 else 0.0075,
                            20.0, 1.0, want_cloud=True)
        d, c = orc.render(s.tris, s.tris_model_count, poses, pmm, W, H, case.proj, src, s.source_label if dof6 else None, lab, 1.0)
        r = orc.costs(d, c, W, H, case.cx, case.cy, case.fx, case.fy, 100.0, stride, lab, obs["aos"], obs["rgb"],
                      obs["label"] if dof6 else None, tot, 2 if dof6 else 0, True, 0.01 if dof6 else 0.0075, 20.0)
        print(f"== dof6={dof6} source={src_name}: n_rend ref {q['n_rend']} oracle {int(r['counts'][0].sum())}")
        print(" rendered ref   ", np.round(q["rendered_cost"], 3))
        print(" rendered oracle", np.round(r["rendered_cost"], 3))
        print(" observed ref   ", np.round(q["observed_cost"], 3))
        print(" observed oracle", np.round(r["observed_cost"], 3))
        oc = orc.depth2cloud(d, c, W, H, len(poses), case.cx, case.cy, case.fx, case.fy, 100.0, stride, pose_label=lab)
        if oc["n"] == q["n_rend"]:
            print(" cloud equal:", np.array_equal(oc["soa"].view(np.uint32), q["cloud"].view(np.uint32)))
            if dof6:
                order = np.argsort(obs["label"], kind="stable")
                tgt = obs["aos"][order]
                sl = np.sort(obs["label"])
                nl = sl.max() + 1
                lb = np.array([np.searchsorted(sl, l, "left") for l in range(nl)], np.int32)
                le = np.array([np.searchsorted(sl, l, "right") for l in range(nl)], np.int32)
                dd, ii = orc.knn_k1(oc["aos"], tgt, oc["label"], lb, le)
            else:
                dd, ii = orc.knn_k1(oc["aos"], obs["aos"])
            print(" nn dist equal:", np.array_equal(dd.view(np.uint32), q["nn_sqdist"].view(np.uint32)), " idx equal:", np.array_equal(ii, q["nn_index"]),
                  " n diff", np.count_nonzero(ii != q["nn_index"]))
            bad = np.nonzero(ii != q["nn_index"])[0][:5]
            for b in bad:
                print("   pt", b, "pose", oc["pose_map"][b], "label", oc["label"][b], "oracle", ii[b], dd[b], "ref", q["nn_index"][b], q["nn_sqdist"][b])

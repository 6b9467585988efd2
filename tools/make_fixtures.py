#!/usr/bin/env python3
"""Extract the reference-embedded known-answer fixtures into tests/golden/.

Run in the build container only (reads /root/reference, which does not exist on the GPU box):
    python tools/make_fixtures.py
Sources:
  B.1  test_develop/test_rnea_derivatives.ipynb cell 7  -> FK oMi[7] at one q
  B.2  test_develop/test_rnea_derivatives.ipynb cell 5  -> world-aligned joint-7 Jacobian * 1
  B.3  analysis/data_analysis.ipynb cell 1 (56 plotly traces: MPC + Ruckig q,v,a,tau, 201 samples)
       analysis/data_analysis.ipynb cell 3 (FK of frame panda_link8 along the MPC positions)
"""
import json, os, sys
import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def main():
    os.makedirs(OUT, exist_ok=True)
    nb = json.load(open(f"{REF}/analysis/data_analysis.ipynb"))
    pj = nb["cells"][1]["outputs"][0]["data"]["application/vnd.plotly.v1+json"]
    tr = pj["data"]
    assert len(tr) == 56
    sig = {}
    for who, off in (("mpc", 0), ("rk", 1)):
        t = np.array(tr[off]["x"], dtype=float)
        arr = np.zeros((4, 7, 201))
        for i in range(7):
            for j in range(4):
                arr[j, i] = np.array(tr[8 * i + 2 * j + off]["y"], dtype=float)
        sig[who] = (t, arr)
    shapes = pj["layout"]["shapes"]
    # hlines: per joint: pos target, pos lim lo, hi, vel target, vel lim +,-, acc +,-, tau +,- = 10 per joint
    ys = [s["y0"] for s in shapes]
    tgt_q = np.array([ys[10 * i + 0] for i in range(7)])
    tgt_v = np.array([ys[10 * i + 3] for i in range(7)])
    fk = nb["cells"][3]["outputs"][0]["data"]["application/vnd.plotly.v1+json"]["data"]
    link8 = np.array([np.array(f["y"], dtype=float) for f in fk])
    np.savez_compressed(
        os.path.join(OUT, "embedded_trajectory.npz"),
        t_mpc=sig["mpc"][0], q_mpc=sig["mpc"][1][0], v_mpc=sig["mpc"][1][1], a_mpc=sig["mpc"][1][2], tau_mpc=sig["mpc"][1][3],
        t_rk=sig["rk"][0], q_rk=sig["rk"][1][0], v_rk=sig["rk"][1][1], a_rk=sig["rk"][1][2], tau_rk=sig["rk"][1][3],
        target_q=tgt_q, target_v=tgt_v, link8_xyz_mpc=link8,
        title=np.array(pj["layout"]["title"]["text"]),
    )
    kat = {
        "fk_oMi7": {
            "q": [0.2, 0.3, -0.1, -1.1, 0.2, 1.4, 0.5],
            "R": [[0.934675, -0.354976, -0.019329], [-0.347142, -0.923067, 0.165649], [-0.0766435, -0.148118, -0.985995]],
            "p": [0.618923, 0.0804767, 0.757975],
            "source": "test_develop/test_rnea_derivatives.ipynb cell 7 output",
        },
        "jac_joint7_world_aligned_times_ones": {
            "q": [0.0, 0.0, 0.0, -1.5, 0.0, 1.0, 0.0],
            "Jqd": [0.35818945, 1.12140565, -0.00527273, 0.51806945, -1.0, 1.19315464],
            "source": "test_develop/test_rnea_derivatives.ipynb cell 5 output",
        },
    }
    json.dump(kat, open(os.path.join(OUT, "kat.json"), "w"), indent=1)
    print("T_rk", sig["rk"][0][-1], "T_mpc", sig["mpc"][0][-1], "target q", tgt_q, "target v", tgt_v)


if __name__ == "__main__":
    sys.exit(main())

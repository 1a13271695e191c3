#!/usr/bin/env python
"""TEST INFRASTRUCTURE.  Regenerates tests/golden/* from the reference-derived probes in oracle/_ref
(which are compiled from sources under /root/reference by oracle/Makefile).  Run in the authoring
container only; the GPU box uses the committed fixtures."""
import json, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "..", "tests", "golden")

def main():
    os.makedirs(GOLD, exist_ok=True)
    subprocess.check_call(["make", "-C", HERE, "ref"])
    out = subprocess.check_output([os.path.join(HERE, "_ref", "enoki_probe")])
    data = json.loads(out)
    with open(os.path.join(GOLD, "enoki_pcg32.json"), "w") as f:
        json.dump(data, f, indent=0)
    print("wrote enoki_pcg32.json:", {k: len(v) for k, v in data.items()})
    probe = os.path.join(HERE, "_ref", "refprobe")
    if os.path.exists(probe):
        out = subprocess.check_output([probe])
        data = json.loads(out)
        with open(os.path.join(GOLD, "ref_vrl_struct.json"), "w") as f:
            json.dump(data, f)
        print("wrote ref_vrl_struct.json:", {k: (len(v) if hasattr(v, "__len__") else v) for k, v in data.items()})
    probe = os.path.join(HERE, "_ref", "refprobe_pm")
    if os.path.exists(probe):
        data = json.loads(subprocess.check_output([probe]))
        with open(os.path.join(GOLD, "ref_photonmap.json"), "w") as f:
            json.dump(data, f)
        print("wrote ref_photonmap.json:", {k: (len(v) if hasattr(v, "__len__") else v) for k, v in data.items()})
    probe = os.path.join(HERE, "_ref", "refprobe_fresnel")
    if os.path.exists(probe):
        data = json.loads(subprocess.check_output([probe]))
        with open(os.path.join(GOLD, "ref_fresnel.json"), "w") as f:
            json.dump(data, f)
        print("wrote ref_fresnel.json:", {k: len(v) for k, v in data.items()})
    probe = os.path.join(HERE, "_ref", "refprobe_nl")
    if os.path.exists(probe):
        data = json.loads(subprocess.check_output([probe]))
        with open(os.path.join(GOLD, "ref_nonlinear.json"), "w") as f:
            json.dump(data, f)
        print("wrote ref_nonlinear.json:", {m["name"]: len(m["rays"]) for m in data["media"]}, "warnings", data["warnings"])
    probe = os.path.join(HERE, "_ref", "refprobe_medium")
    if os.path.exists(probe):
        data = json.loads(subprocess.check_output([probe]))
        with open(os.path.join(GOLD, "ref_medium.json"), "w") as f:
            json.dump(data, f)
        print("wrote ref_medium.json:", len(data["cases"]), "cases, warnings", data["warnings"])
    probe = os.path.join(HERE, "_ref", "refprobe_lc")
    if os.path.exists(probe):
        data = json.loads(subprocess.check_output([probe]))
        with open(os.path.join(GOLD, "ref_lightcut.json"), "w") as f:
            json.dump(data, f)
        print("wrote ref_lightcut.json:", {k: (len(v) if hasattr(v, "__len__") else v) for k, v in data.items()})

if __name__ == "__main__":
    sys.exit(main())

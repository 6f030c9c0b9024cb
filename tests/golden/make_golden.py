"""Generates the committed fixtures from the reference's own test data (run in the build container only;
/root/reference does not exist on the GPU box).

  scenes/<name>/         verbatim copies of the reference's test scenes (tests/scenes/*: scene.isf + PNG textures):
                         golden INPUTS shared by the oracle and the CUDA path.
  moller_trumbore.npz    the 3012 hit + 3012 miss vectors of tests/moller_trumbore/{hit,miss}_tests.yml as float64
                         arrays (ray o/d, triangle v0 v1 v2, gold dist/u/v) — the known answers of
                         src/scene/internal/triangle.rs:200-233.
  sha1_kat.json          the seven image hashes asserted by src/main.rs:100-165 with their profiles.
"""
import json
import os
import shutil

import numpy as np
import yaml

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    dst = os.path.join(HERE, "scenes")
    for name in sorted(os.listdir(os.path.join(REF, "tests/scenes"))):
        shutil.copytree(os.path.join(REF, "tests/scenes", name), os.path.join(dst, name), dirs_exist_ok=True)
    out = {}
    for kind in ("hit", "miss"):
        with open(os.path.join(REF, "tests/moller_trumbore/%s_tests.yml" % kind)) as f:
            tests = yaml.load(f, Loader=yaml.CSafeLoader)
        rays = np.array([t["ray"]["position"] + t["ray"]["direction"] for t in tests], np.float64)
        tris = np.array([t["triangle"]["v0"] + t["triangle"]["v1"] + t["triangle"]["v2"] for t in tests], np.float64)
        out[kind + "_rays"] = rays
        out[kind + "_tris"] = tris
        if kind == "hit":
            out["hit_gold"] = np.array([[t["hit"]["dist"], t["hit"]["u"], t["hit"]["v"]] for t in tests], np.float64)
    np.savez_compressed(os.path.join(HERE, "moller_trumbore.npz"), **out)
    kat = {
        "profile": {"width": 800, "height": 600, "samples": 16, "bounces": 4, "tonemap": "FILMIC"},
        "scenes": {
            "cube": "60558456ace7e8063ebfab219ee35a2c7de862f5",
            "reflection": "6ccc3b9f20442f15f25c41cf8d342ede5185e3db",
            "head": "2c90976144ba14fe9f06ec3c812ff30f0a0c9146",
            "spheres": "fe2687e274ac978a4815f202612eca71ee8dd8c9",
            "alpha_transparency": "fdf9ccbe9dc3f3102e3c05b96d2984000e73b62f",
            "white_furnace_indirect": "80dd0598ced75660b80170e69cad1a74fba26a15",
            "white_furnace_direct": "6838e727798bd33f2f796be3edaa893445087159",
        },
        "bounces_override": {"white_furnace_direct": 0},
        "source": "src/main.rs:100-165",
    }
    with open(os.path.join(HERE, "sha1_kat.json"), "w") as f:
        json.dump(kat, f, indent=1)


if __name__ == "__main__":
    main()

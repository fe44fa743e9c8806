"""Generates tests/golden/mbm_<robot>.json.gz from the reference checkout (run in the CPU container).

Follows resources/problem_tar_to_pkl_json.py:28-76,93-127 of the reference: every MotionBenchMaker
(scene, request) YAML pair of resources/<robot>/problems.tar.bz2 becomes one problem dict
{sphere, cylinder, box, start, goals, index, problem}.  The reference's own pure-Python
`vamp/transformations.py` is imported from the checkout for the pose algebra (it is not copied).
The `valid` label of the reference (validate(start) and any(validate(goal))) is produced separately
by the oracle in tests, since the reference's Python module is not built here.

usage: python tests/golden/make_mbm_fixture.py [--ref /root/reference] [--keep N] robot...
"""
import argparse
import gzip
import importlib.util
import json
import os
import re
import tarfile
from collections import defaultdict

import numpy as np
import yaml

try:
    from yaml import CLoader as Loader
except ImportError:
    from yaml import Loader

HERE = os.path.dirname(os.path.abspath(__file__))

# src/vamp/constants.py ROBOT_JOINTS
ROBOT_JOINTS = {
    "panda": [f"panda_joint{i}" for i in range(1, 8)],
    "ur5": ["shoulder_pan_joint", "shoulder_lift_joint", "elbow_joint", "wrist_1_joint", "wrist_2_joint", "wrist_3_joint"],
    "fetch": ["torso_lift_joint", "shoulder_pan_joint", "shoulder_lift_joint", "upperarm_roll_joint", "elbow_flex_joint",
              "forearm_roll_joint", "wrist_flex_joint", "wrist_roll_joint"],
    "baxter": ["right_s0", "right_s1", "right_e0", "right_e1", "right_w0", "right_w1", "right_w2",
               "left_s0", "left_s1", "left_e0", "left_e1", "left_w0", "left_w1", "left_w2"],
}


def load_transformations(ref):
    spec = importlib.util.spec_from_file_location("vamp_transformations", os.path.join(ref, "src/vamp/transformations.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("--keep", type=int, default=0, help="keep only the first N problems of each scenario (0 = all)")
    ap.add_argument("robots", nargs="+")
    args = ap.parse_args()
    T = load_transformations(args.ref)

    def tf(t):
        return np.dot(T.translation_matrix(t["position"]), T.quaternion_matrix(t["orientation"]))

    def scene(data):
        objects = {"sphere": [], "cylinder": [], "box": []}
        for co in data["world"]["collision_objects"]:
            obj = {"name": co["id"]}
            base = T.identity_matrix() if "pose" not in co else tf(co["pose"])
            prim, pose = co["primitives"][0], np.dot(base, tf(co["primitive_poses"][0]))
            kind = prim["type"]
            if kind == "sphere":
                obj["radius"] = prim["dimensions"][0]
            elif kind == "cylinder":
                obj["length"], obj["radius"] = prim["dimensions"][0], prim["dimensions"][1]
            elif kind == "box":
                obj["half_extents"] = [x / 2 for x in prim["dimensions"]]
            else:
                raise RuntimeError(kind)
            obj["position"] = T.translation_from_matrix(pose).tolist()
            obj["orientation_euler_xyz"] = [float(v) for v in T.euler_from_matrix(pose)]
            objects[kind].append(obj)
        return objects

    def request(data, joints):
        js = data["start_state"]["joint_state"]
        start = [js["position"][js["name"].index(j)] for j in joints]
        jc = data["goal_constraints"][0]["joint_constraints"]
        names, pos = [e["joint_name"] for e in jc], [e["position"] for e in jc]
        return {"start": start, "goals": [[pos[names.index(j)] for j in joints]]}

    for robot in args.robots:
        joints = ROBOT_JOINTS[robot]
        scenes, requests = defaultdict(dict), defaultdict(dict)
        tar = tarfile.open(os.path.join(args.ref, "resources", robot, "problems.tar.bz2"), "r:bz2")
        for m in tar.getmembers():
            if not m.isfile():
                continue
            _, problem, filename = m.name.split("/")
            problem = problem.replace(f"_{robot}", "")
            index = int(re.findall(r"\d+", filename)[0])
            if args.keep and index > args.keep:
                continue
            data = yaml.load(tar.extractfile(m).read(), Loader=Loader)
            if "scene" in filename:
                scenes[problem][index] = scene(data)
            elif "request" in filename:
                requests[problem][index] = request(data, joints)
        out = {"robot": robot, "joints": joints, "problems": {}}
        for k in sorted(scenes):
            out["problems"][k] = [
                {**scenes[k][i], **requests[k][i], "index": i, "problem": k} for i in sorted(scenes[k]) if i in requests[k]
            ]
        path = os.path.join(HERE, f"mbm_{robot}.json.gz")
        with gzip.open(path, "wt") as fh:
            json.dump(out, fh, separators=(",", ":"))
        print(robot, {k: len(v) for k, v in out["problems"].items()}, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

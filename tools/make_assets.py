#!/usr/bin/env python3
"""Normalise the reference's *input data* (MJCF models, task YAMLs, TestTasks CSV fixtures)
into assets/ so that tests / bench can run on a box where /root/reference does not exist.

Only dynamics-relevant content is kept: <include>s are resolved, visual-only elements
(textures, materials, skybox, cameras, lights, sensors, hfield assets) and visual attributes
(material, rgba, group) are dropped, comments are dropped.  Nothing here is reference *source*
code; these are the inputs both the reference and this implementation consume.

Sources (relative to /root/reference):
  mujoco_models/Acrobot/acrobot.xml            -> assets/models/acrobot.xml
  mujoco_models/walker/walker_plane.xml        -> assets/models/walker_plane.xml
  TaskConfigs/toys/acrobot.yaml                -> assets/tasks/acrobot.yaml
  TaskConfigs/locomotion/run_plane.yaml        -> assets/tasks/walker_run.yaml
  TestTasks/{acrobot,walker_run}/{0..99}.csv   -> assets/test_tasks/<task>.csv (one row per trial)

Run:  python tools/make_assets.py [/root/reference]
"""
import os
import sys
import xml.etree.ElementTree as ET

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "assets")

DROP_TAGS = {"texture", "material", "camera", "light", "sensor", "visual", "statistic", "hfield", "size",
             "skin", "keyframe"}
DROP_ATTRS = {"material", "rgba", "group"}
VISUAL_INCLUDES = ("visual.xml", "skybox.xml", "materials.xml")


def resolve_includes(elem, base_dir):
    out = []
    for child in list(elem):
        if child.tag == "include":
            f = child.get("file")
            if f.endswith(VISUAL_INCLUDES):
                elem.remove(child)
                continue
            sub = ET.parse(os.path.join(base_dir, f)).getroot()
            resolve_includes(sub, os.path.dirname(os.path.join(base_dir, f)))
            idx = list(elem).index(child)
            elem.remove(child)
            # an included file's root is <mujoco> (or <mujocoinclude>): splice its children
            for k, c in enumerate(list(sub)):
                elem.insert(idx + k, c)
        else:
            resolve_includes(child, base_dir)
    return out


def strip(elem):
    for child in list(elem):
        if child.tag in DROP_TAGS:
            elem.remove(child)
            continue
        strip(child)
    for a in list(elem.attrib):
        if a in DROP_ATTRS:
            del elem.attrib[a]
    elem.text = None
    elem.tail = None


def prune_empty(elem):
    for child in list(elem):
        prune_empty(child)
        if child.tag in ("asset", "default") and len(child) == 0 and child.tag == "asset":
            elem.remove(child)


def indent(elem, level=0):
    pad = "\n" + "  " * level
    if len(elem):
        elem.text = pad + "  "
        for c in elem:
            indent(c, level + 1)
            c.tail = pad + "  "
        elem[-1].tail = pad
    if level == 0:
        elem.tail = "\n"


def convert_model(src, dst, header):
    root = ET.parse(src).getroot()
    resolve_includes(root, os.path.dirname(src))
    strip(root)
    prune_empty(root)
    indent(root)
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    with open(dst, "w") as f:
        f.write("<!-- %s -->\n" % header)
        f.write(ET.tostring(root, encoding="unicode"))


def needed_meshes(root):
    """Mesh names used by geoms that take part in the dynamics (they collide or carry mass).  Visual-only geoms
    (contype = conaffinity = 0 and mass = 0) keep their place in the geom list but their mesh files are not shipped."""
    classes = {"main": {}}

    def walk_defaults(node, parent):
        name = node.get("class", "main")
        attrs = dict(classes.get(name, {})) if name == "main" else dict(parent)
        for ch in node:
            if ch.tag == "geom":
                attrs.update(ch.attrib)
        classes[name] = attrs
        for ch in node:
            if ch.tag == "default":
                walk_defaults(ch, attrs)

    for d in root.findall("default"):
        walk_defaults(d, classes["main"])
    used = set()

    def walk_body(node, childclass):
        childclass = node.get("childclass", childclass)
        for ch in node:
            if ch.tag == "geom":
                a = dict(classes.get(ch.get("class", childclass or "main"), {}))
                a.update(ch.attrib)
                if "mesh" not in a:
                    continue
                collides = int(a.get("contype", 1)) != 0 or int(a.get("conaffinity", 1)) != 0
                massive = float(a.get("mass", 1)) != 0
                if collides or massive:
                    used.add(a["mesh"])
            elif ch.tag == "body":
                walk_body(ch, childclass)

    for wb in root.findall("worldbody"):
        walk_body(wb, None)
    return used


def convert_franka_scene(src, dst, header):
    """Franka scenes: as convert_model, plus the mesh files the dynamics needs (copied next to the model, keeping
    the paths the <mesh file=...> attributes name; unused <mesh> assets stay listed but are never opened)."""
    import shutil
    root = ET.parse(src).getroot()
    resolve_includes(root, os.path.dirname(src))
    strip(root)
    prune_empty(root)
    used = needed_meshes(root)
    for asset in root.findall("asset"):
        for mesh in asset.findall("mesh"):
            name = mesh.get("name") or os.path.splitext(os.path.basename(mesh.get("file")))[0]
            if name in used:
                f = mesh.get("file")
                out = os.path.join(os.path.dirname(dst), f)
                os.makedirs(os.path.dirname(out), exist_ok=True)
                shutil.copyfile(os.path.join(os.path.dirname(src), f), out)
                os.chmod(out, 0o644)
    indent(root)
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    with open(dst, "w") as f:
        f.write("<!-- %s -->\n" % header)
        f.write(ET.tostring(root, encoding="unicode"))


def copy_yaml(src, dst):
    import yaml
    with open(src) as f:
        cfg = yaml.safe_load(f)
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    with open(dst, "w") as f:
        f.write("# normalised from %s by tools/make_assets.py\n" % os.path.relpath(src, REF))
        yaml.safe_dump(cfg, f, sort_keys=False, default_flow_style=None)


def pack_test_tasks(task, dst):
    rows = []
    for i in range(100):
        p = os.path.join(REF, "TestTasks", task, "%d.csv" % i)
        if not os.path.exists(p):
            break
        with open(p) as f:
            line = f.readline().strip().rstrip(",")
        rows.append(line)
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    with open(dst, "w") as f:
        f.write("# one TestTasks/%s/<row>.csv per line (start state then residual targets)\n" % task)
        for r in rows:
            f.write(r + "\n")


def main():
    mm = os.path.join(REF, "mujoco_models")
    convert_model(os.path.join(mm, "Acrobot/acrobot.xml"), os.path.join(OUT, "models/acrobot.xml"),
                  "dynamics-only normalisation of mujoco_models/Acrobot/acrobot.xml (tools/make_assets.py)")
    convert_model(os.path.join(mm, "walker/walker_plane.xml"), os.path.join(OUT, "models/walker_plane.xml"),
                  "dynamics-only normalisation of mujoco_models/walker/walker_plane.xml (tools/make_assets.py)")
    copy_yaml(os.path.join(REF, "TaskConfigs/toys/acrobot.yaml"), os.path.join(OUT, "tasks/acrobot.yaml"))
    copy_yaml(os.path.join(REF, "TaskConfigs/locomotion/run_plane.yaml"), os.path.join(OUT, "tasks/walker_run.yaml"))
    fr = os.path.join(mm, "Franka_emika_scenes_V1")
    for scene in ("reaching_scene", "boxSweep", "cylinder_pushing", "cylinder_pushing_lowClutter",
                  "cylinder_pushing_heavyClutter"):
        convert_franka_scene(os.path.join(fr, scene + ".xml"), os.path.join(OUT, "models/franka/%s.xml" % scene),
                             "dynamics-only normalisation of mujoco_models/Franka_emika_scenes_V1/%s.xml "
                             "(tools/make_assets.py)" % scene)
    for src, name in (("free_motion/reaching.yaml", "reaching"), ("rigid_body_manipulation/box_sweep.yaml", "box_sweep"),
                      ("rigid_body_manipulation/twoD_push_no_clutter.yaml", "push_ncl"),
                      ("rigid_body_manipulation/twoD_push_light_clutter.yaml", "push_lcl"),
                      ("rigid_body_manipulation/twoD_push_heavy_clutter.yaml", "push_mcl")):
        copy_yaml(os.path.join(REF, "TaskConfigs", src), os.path.join(OUT, "tasks/%s.yaml" % name))
    for task in ("reaching", "box_sweep", "push_ncl", "push_lcl", "push_mcl"):
        pack_test_tasks(task, os.path.join(OUT, "test_tasks/%s.csv" % task))
    pack_test_tasks("acrobot", os.path.join(OUT, "test_tasks/acrobot.csv"))
    pack_test_tasks("walker_run", os.path.join(OUT, "test_tasks/walker_run.csv"))
    print("assets written to", os.path.normpath(OUT))


if __name__ == "__main__":
    main()

"""Pins the Cornell fixture to the REFERENCE's own source: the literals, expressions, triangle order,
winding and material assignment are read out of /root/reference/RTrace/scene.swift:14-240 with regular expressions,
evaluated in fp32 the way Swift's Float would, and compared with the bytes of the two restatements this repo ships
(oracle/rt_fixtures.c, csrc/host_scene.cpp).  Exact for every value that is a literal or a sum of literals; the rotated
box corners go through cos/sin and a 4x4 product whose rounding Apple's simd does not pin: <= 2 ulp there.
Skipped where the reference is absent (the GPU box)."""
import os
import re

import numpy as np
import pytest

import oracle_lib as O

SCENE = "/root/reference/RTrace/scene.swift"
pytestmark = pytest.mark.skipif(not os.path.exists(SCENE), reason="reference sources not present")
F = np.float32


def fp32_eval(expr, env):
    """A Swift Float expression of scene.swift: literals, names, + - * /, Float.pi, unary minus -- every operand fp32."""
    expr = expr.replace("Float.pi", "PI")
    expr = re.sub(r"(?<![\w.])(\d+\.\d+|\d+)(?![\w.])", r"F(\1)", expr)
    return F(eval(expr, {"F": F, "PI": F(np.pi), "__builtins__": {}}, env))


def body(src, name):
    start = src.index(f"func {name}(")
    depth, i = 0, src.index("{", start)
    for j in range(i, len(src)):
        depth += src[j] == "{"
        depth -= src[j] == "}"
        if depth == 0:
            return src[i:j]
    raise AssertionError(name)


def strip_comments(s):
    return re.sub(r"//[^\n]*", "", s)


def vec3(text, env):
    return [fp32_eval(p, env) for p in text.split(",")]


def scene_from_swift():
    src = open(SCENE).read()
    room, init = strip_comments(body(src, "createCornellBoxScene")), strip_comments(body(src, "initCornellBox"))
    rot, boxt = strip_comments(body(src, "createRotatedBoxVertices")), strip_comments(body(src, "createBoxTriangles"))
    env = {}
    for name, expr in re.findall(r"let (\w+)(?:: Float)? = ([^\n(]+?)\n", room):          # roomSize, half, box sizes, rotations
        env[name] = fp32_eval(expr, env)
    mats = {m[0]: (vec3(m[1], env), F(m[2]), F(m[3])) for m in
            re.findall(r"let (\w+) = Material\(diffuse: simd_float4\(([^)]*)\), metallic: ([\d.]+), roughness: ([\d.]+)\)", room)}
    tris = []                                                                               # (3 vertices, material name)
    for verts, mat in re.findall(r"Triangle\(vertices: \[(.*?)\], material: (\w+)\)", room, re.S):
        tris.append(([vec3(v, env) for v in re.findall(r"simd_float3\(([^)]*)\)", verts)], mat))
    # the two boxes: position, size, rotation -> createRotatedBoxVertices -> createBoxTriangles
    signs = [[(-1 if s else 1) for s in (a == "-", b == "-", c == "-")]
             for a, b, c in re.findall(r"simd_float3\(\s*(-?)halfW,\s*(-?)halfH,\s*(-?)halfD\)", rot)]
    cols = [c.replace(" ", "").split(",") for c in re.findall(r"simd_float4\(([^)]*)\)", rot.split("rotationMatrix = simd_float4x4(")[1].split("\n    )")[0])]
    order = [tuple(int(i) for i in t) for t in re.findall(r"vertices\[(\d)\], vertices\[(\d)\], vertices\[(\d)\]", boxt)]
    assert len(signs) == 8 and len(cols) == 4 and len(order) == 12
    for pos, call in re.findall(r"let \w+Position = simd_float3\(([^)]*)\).*?createRotatedBoxVertices\((.*?)\)", room, re.S):
        args = dict(re.findall(r"(\w+): (\w+)", call))
        c = vec3(pos, env)
        hw, hh, hd = env[args["width"]] / F(2), env[args["height"]] / F(2), env[args["depth"]] / F(2)
        cy, sy = F(np.cos(env[args["rotationY"]])), F(np.sin(env[args["rotationY"]]))
        sym = {"cosY": cy, "sinY": sy, "-sinY": -sy, "0": F(0), "1": F(1)}
        M = np.array([[sym[e] for e in col] for col in cols], F).T                         # simd_float4x4 takes COLUMNS
        corners = []
        for sx, sh, sd in signs:
            v = np.array([F(sx) * hw, F(sh) * hh, F(sd) * hd, F(1)], F)
            r = [F(sum(float(M[i, k]) * float(v[k]) for k in range(4))) for i in range(3)]  # one rounding: the reference's is unpinned
            corners.append([F(r[0] + c[0]), F(r[1] + c[1]), F(r[2] + c[2])])
        material = re.search(r"\n\s*triangles\.append\(contentsOf: createBoxTriangles\(vertices: \w+, material: (\w+)\)\)", room).group(1)
        for a, b, d in order:
            tris.append(([corners[a], corners[b], corners[d]], material))
    # the light: initCornellBox
    lenv = {}
    for name, expr in re.findall(r"let (\w+)(?:: Float)? = ([^\n(]+?)\n", init):
        lenv[name] = fp32_eval(expr, lenv)
    center = vec3(re.search(r"let lightCenter = simd_float3\(([^)]*)\)", init).group(1), lenv)
    lenv.update({"lightCenter_x": center[0], "lightCenter_z": center[2]})
    lv = {n: vec3(e.replace("lightCenter.x", "lightCenter_x").replace("lightCenter.z", "lightCenter_z"), lenv)
          for n, e in re.findall(r"let (v\d) = simd_float3\(([^)]*)\)", init)}
    lm = re.search(r"Material\(\s*diffuse: simd_float4\(([^)]*)\),\s*metallic: ([\d.]+),\s*roughness: ([\d.]+),\s*emissive: simd_float3\(([^)]*)\)", init)
    mats["lightMaterial"] = (vec3(lm.group(1), {}), F(lm.group(2)), F(lm.group(3)), vec3(lm.group(4), {}))
    for names in re.findall(r"Triangle\(vertices: \[(v\d), (v\d), (v\d)\], material: lightMaterial\)", init):
        tris.append(([lv[n] for n in names], "lightMaterial"))
    cam = re.search(r"simd_normalize\(simd_float3\(([^)]*)\) - simd_float3\(([^)]*)\)\)", init)
    camera = dict(target=vec3(cam.group(1), {}), position=vec3(cam.group(2), {}),
                  fov=fp32_eval(re.search(r"horizontalFov: ([^)]*)\)", init).group(1), {}))
    light = dict(center=center, width=lenv["lightWidth"], depth=lenv["lightDepth"])
    return tris, mats, camera, light


def ulps(a, b):
    a, b = np.asarray(a, F).view(np.int32).astype(np.int64), np.asarray(b, F).view(np.int32).astype(np.int64)
    return int(np.abs(a - b).max())


def test_fixture_bytes_are_scene_swift():
    tris, mats, camera, light = scene_from_swift()
    assert len(tris) == 36
    sc = O.cornell(800, 600)
    V = sc["vertices"].view(F).reshape(36, 3, 4)
    M = sc["materials"].view(F).reshape(36, 12)
    for t, (verts, mat) in enumerate(tris):
        want = np.array(verts, F)
        exact = t < 10 or t >= 34                          # walls and the light: literals and sums of literals
        assert ulps(V[t, :, :3], want) <= (0 if exact else 2), (t, V[t, :, :3], want)
        m = mats[mat]
        assert np.array_equal(M[t, 0:4], np.array(m[0], F)) and M[t, 4] == m[1] and M[t, 5] == m[2], (t, mat)
        assert np.array_equal(M[t, 8:11], np.array(m[3] if len(m) > 3 else [0, 0, 0], F)), (t, mat)
    # shared corners stay shared bit for bit in the fixture (what the object groups of the sorted kernel rest on)
    assert np.array_equal(V[10, 0], V[11, 0]) and np.array_equal(V[22, 0], V[23, 0])
    cam = sc["camera"].view(F)
    assert np.array_equal(cam[0:3], np.array(camera["position"], F))
    d = np.array(camera["target"], F) - np.array(camera["position"], F)
    assert np.array_equal(cam[4:7], (d / F(np.sqrt((d * d).sum()))).astype(F))
    assert cam[14] == camera["fov"]
    L = sc["lights"].view(F)
    assert np.array_equal(L[0:3], np.array(light["center"], F)) and L[12] == light["width"] and L[13] == light["depth"]
    assert np.array_equal(L[4:8], np.array(mats["lightMaterial"][0], F))      # convertSquareLight: color = material.diffuse


def test_host_mirror_is_the_same_bytes():
    """csrc/host_scene.cpp (the product's scene mirror) against the fixture just pinned: byte for byte."""
    import gpuraytracer_b200 as G
    g, o = G.init_cornell_box(800, 600), O.cornell(800, 600)
    for key in ("camera", "materials", "lights", "vertices"):
        assert np.array_equal(getattr(g, key), o[key]), key

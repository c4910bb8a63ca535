"""Regenerates this repo's own fixtures under ray-tracer_b200/assets/ (SURVEY.md 8(f) rows 1 and 3).

  textures/earthlike.png, textures/checker.png   synthetic images (the reference ships none: its
                                                  scene 2 wants an `earth.png` that is not in the repo)
  textures/parsed_textures.{txt,bin}             written by ray-tracer_b200/textures.py
  models/studio.obj, models/studio.mtl           four material groups: image-textured crate (vt),
                                                  glass gem (d < 1, Ni), emissive panel (Ke), glossy pyramid (Ns)
  tests/golden/parsed_textures_reference.txt     written by the REFERENCE's own textures/parse_textures.py
                                                  run on the same directory (only where /root/reference exists)
Deterministic: running it again reproduces the committed files.
"""
import os
import shutil
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ASSETS = os.path.join(ROOT, "ray-tracer_b200", "assets")
REF_SCRIPT = "/root/reference/textures/parse_textures.py"


def earthlike(w=64, h=32):
    """blue 'ocean', green/brown 'continents' from a few fixed sinusoids, white caps: any smooth
    image with all three channels varying would do"""
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    lon, lat = x / w * 2 * np.pi, (y / (h - 1) - 0.5) * np.pi
    f = np.sin(3 * lon + 1.0) * np.cos(2 * lat) + 0.6 * np.sin(5 * lon - 2 * lat + 0.5) + 0.4 * np.cos(7 * lat + lon)
    land = f > 0.25
    img = np.zeros((h, w, 3), np.float64)
    img[..., 0] = np.where(land, 90 + 60 * np.clip(f, 0, 1), 20 + 10 * np.cos(lat))
    img[..., 1] = np.where(land, 140 - 40 * np.clip(f - 0.5, 0, 1), 60 + 40 * np.cos(lat))
    img[..., 2] = np.where(land, 50, 150 + 80 * np.cos(lat) ** 2)
    caps = np.abs(lat) > 1.25
    img[caps] = 240
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def checker(n=8):
    y, x = np.mgrid[0:n, 0:n]
    v = np.where((x + y) % 2 == 0, 255, 32).astype(np.uint8)
    return np.stack([v, np.roll(v, 1, axis=1), np.full_like(v, 128)], axis=-1)


def box(cx, cy, cz, sx, sy, sz):
    return [(cx + dx * sx, cy + dy * sy, cz + dz * sz) for dx in (-0.5, 0.5) for dy in (-0.5, 0.5) for dz in (-0.5, 0.5)]


def write_models():
    v, vt, lines = [], [], []

    def add(points):
        v.extend(points)
        return list(range(len(v) - len(points) + 1, len(v) + 1))

    # crate: a box of quads, every face mapped to the whole image
    vt.extend([(0.0, 0.0), (1.0, 0.0), (1.0, 1.0), (0.0, 1.0)])
    c = add(box(-0.45, -0.55, 0.1, 0.9, 0.9, 0.9))       # index = 4*ix + 2*iy + iz
    crate_faces = [(c[0], c[1], c[3], c[2]), (c[4], c[6], c[7], c[5]), (c[0], c[4], c[5], c[1]),
                   (c[2], c[3], c[7], c[6]), (c[0], c[2], c[6], c[4]), (c[1], c[5], c[7], c[3])]
    lines.append("usemtl crate")
    lines += ["f " + " ".join(f"{a}/{k + 1}" for k, a in enumerate(face)) for face in crate_faces]

    # gem: an octahedron of 8 triangles, glass
    gx, gy, gz, r = 0.75, -0.45, -0.55, 0.5
    g = add([(gx + r, gy, gz), (gx - r, gy, gz), (gx, gy + 1.2 * r, gz), (gx, gy - 1.1 * r, gz), (gx, gy, gz + r), (gx, gy, gz - r)])
    lines.append("usemtl glass")
    for a, b in ((0, 4), (4, 1), (1, 5), (5, 0)):
        lines.append(f"f {g[a]} {g[b]} {g[2]}")
        lines.append(f"f {g[b]} {g[a]} {g[3]}")

    # panel: a small light, slightly tilted so that no box of the reference's BVH is flat
    p = add([(-0.6, 1.55, -0.4), (0.6, 1.5, -0.4), (0.6, 1.52, 0.6), (-0.6, 1.57, 0.6)])
    lines.append("usemtl panel")
    lines.append(f"f {p[0]} {p[1]} {p[2]} {p[3]}")

    # pyramid: glossy red, checker-free; uses negative (relative) indices for its apex
    lid = add([(-0.9, -0.05, -0.35), (0.0, -0.08, -0.35), (0.0, -0.02, 0.55), (-0.9, -0.06, 0.55)])
    q = add([(0.35, -1.0, 0.35), (1.25, -1.0, 0.45), (1.15, -1.0, 1.3), (0.3, -0.98, 1.2), (0.8, 0.1, 0.8)])
    lines.append("usemtl lacquer")
    for a, b in ((0, 1), (1, 2), (2, 3), (3, 0)):
        lines.append(f"f {q[a]} {q[b]} -1")   # the apex is the last vertex of the file
    # crate again: a second group of the same material later in the file (a lid), must join the first
    lines.append("usemtl crate")
    lines.append("f " + " ".join(f"{a}/{k + 1}" for k, a in enumerate(lid)))

    os.makedirs(os.path.join(ASSETS, "models"), exist_ok=True)
    with open(os.path.join(ASSETS, "models", "studio.obj"), "w") as f:
        f.write("# studio.obj -- fixture of ray-tracer_b200 (tools/make_assets.py)\nmtllib studio.mtl\no studio\n")
        f.write("".join(f"v {x:.6f} {y:.6f} {z:.6f}\n" for x, y, z in v))
        f.write("vn 0.0000 1.0000 0.0000\n")
        f.write("".join(f"vt {a:.6f} {b:.6f}\n" for a, b in vt))
        f.write("s 0\n" + "\n".join(lines) + "\n")
    with open(os.path.join(ASSETS, "models", "studio.mtl"), "w") as f:
        f.write("# studio.mtl -- fixture of ray-tracer_b200 (tools/make_assets.py)\n\n"
                "newmtl crate\nNs 0.000000\nKa 1.000000 1.000000 1.000000\nKd 0.800000 0.800000 0.800000\nKe 0.000000 0.000000 0.000000\n"
                "Ni 1.450000\nd 1.000000\nillum 2\nmap_Kd textures/earthlike.png\n\n"
                "newmtl glass\nNs 900.000000\nKd 0.950000 1.000000 0.950000\nKe 0.000000 0.000000 0.000000\nNi 1.500000\nd 0.200000\nillum 7\n\n"
                "newmtl panel\nKd 0.000000 0.000000 0.000000\nKe 5.000000 4.500000 3.500000\nNi 1.000000\nd 1.000000\nillum 2\n\n"
                "newmtl lacquer\nNs 810.000000\nKd 0.800000 0.150000 0.100000\nKe 0.000000 0.000000 0.000000\nNi 1.450000\nd 1.000000\nillum 2\n")


def main():
    from PIL import Image
    import raytracer_b200  # noqa: F401  (makes the package importable by name)
    from raytracer_b200 import textures
    tdir = os.path.join(ASSETS, "textures")
    os.makedirs(tdir, exist_ok=True)
    for name in os.listdir(tdir):
        os.remove(os.path.join(tdir, name))
    Image.fromarray(earthlike()).save(os.path.join(tdir, "earthlike.png"))
    Image.fromarray(checker()).save(os.path.join(tdir, "checker.png"))
    if os.path.exists(REF_SCRIPT):
        # the reference's script scans and writes in its current directory
        subprocess.run([sys.executable, REF_SCRIPT], cwd=tdir, check=True)
        shutil.move(os.path.join(tdir, "parsed_textures.txt"), os.path.join(ROOT, "tests", "golden", "parsed_textures_reference.txt"))
    textures.parse_directory(tdir)
    write_models()
    print("assets written under", ASSETS)


if __name__ == "__main__":
    main()

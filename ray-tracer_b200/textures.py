"""Texture pre-parser: the offline step of the reference's image-texture path.

The reference converts every image of a directory into ONE text file with a Python script
(`textures/parse_textures.py:29-58`) that its `ImageTexture` class reads back at scene build time
(`src/main.cu:40-91`).  This module keeps that contract byte for byte:

    <count>\\n  then per image:  <file name>\\n<width>\\n<height>\\n<r g b r g b ... each value = x / 256, each followed by a space>\\n

and adds a binary fast path (`parsed_textures.bin`) holding the same float values, which the C++
loader (`ray-tracer_b200/host/obj_mtl.hpp: load_parsed_texture`) prefers when it is present:

    char magic[8] = "RTBTEX1\\0"; u32 count; per image { u32 name_len; name; u32 width; u32 height; f32 rgb[w*h*3] }

`x / 256` (x an 8-bit channel) is exact in binary floating point, so text and binary hold identical
values.  Host-side tooling: nothing here runs on the render path.

    python -m raytracer_b200.textures <directory>      # writes parsed_textures.txt and .bin there
"""
import os
import struct
import sys

import numpy as np

OUTPUT_TEXT = "parsed_textures.txt"
OUTPUT_BINARY = "parsed_textures.bin"
IMAGE_EXTENSIONS = ("png", "jpg", "jpeg")     # textures/parse_textures.py:7
MAGIC = b"RTBTEX1\0"


def read_image(path):
    """(height, width, 3) uint8, RGB, row 0 = top: what `Image.convert("RGB").getpixel((x, y))`
    yields pixel by pixel (textures/parse_textures.py:19-27).  Binary PPM is read natively so that
    the tool also works without Pillow."""
    if path.lower().endswith(".ppm"):
        with open(path, "rb") as f:
            tokens, data = [], f.read()
        pos = 0
        while len(tokens) < 4:
            while data[pos:pos + 1].isspace():
                pos += 1
            if data[pos:pos + 1] == b"#":
                pos = data.index(b"\n", pos) + 1
                continue
            end = pos
            while not data[end:end + 1].isspace():
                end += 1
            tokens.append(data[pos:end])
            pos = end
        if tokens[0] != b"P6" or int(tokens[3]) != 255:
            raise ValueError(f"{path}: only 8-bit binary PPM (P6) is supported")
        w, h = int(tokens[1]), int(tokens[2])
        return np.frombuffer(data, np.uint8, count=w * h * 3, offset=pos + 1).reshape(h, w, 3).copy()
    from PIL import Image  # the reference's own dependency
    return np.asarray(Image.open(path).convert("RGB"), dtype=np.uint8).copy()


def list_images(directory):
    """the directory scan of get_textures() (textures/parse_textures.py:38-49), in listdir order"""
    names = []
    for name in os.listdir(directory):
        parts = name.split(".")
        if len(parts) == 2 and parts[1] in IMAGE_EXTENSIONS:
            names.append(name)
    return names


def text_block(name, pixels):
    """Texture.get_string(): the values are written with Python's float repr of x / 256"""
    h, w, _ = pixels.shape
    table = [f"{x / 256} " for x in range(256)]
    body = "".join(table[v] for v in pixels.reshape(-1).tolist())
    return f"{name}\n{w}\n{h}\n{body}\n"


def write_text(path, images):
    """images: list of (name, uint8 H x W x 3)"""
    with open(path, "w") as f:
        f.write(f"{len(images)}\n" + "".join(text_block(n, p) for n, p in images))


def write_binary(path, images):
    with open(path, "wb") as f:
        f.write(MAGIC + struct.pack("<I", len(images)))
        for name, pixels in images:
            h, w, _ = pixels.shape
            raw = name.encode()
            f.write(struct.pack("<I", len(raw)) + raw + struct.pack("<II", w, h))
            f.write((pixels.astype(np.float32) / np.float32(256)).astype("<f4").tobytes())


def read_binary(path):
    """-> {name: float32 H x W x 3}"""
    out = {}
    with open(path, "rb") as f:
        if f.read(8) != MAGIC:
            raise ValueError(f"{path}: not a parsed-texture binary")
        (count,) = struct.unpack("<I", f.read(4))
        for _ in range(count):
            (n,) = struct.unpack("<I", f.read(4))
            name = f.read(n).decode()
            w, h = struct.unpack("<II", f.read(8))
            out[name] = np.frombuffer(f.read(w * h * 12), "<f4").reshape(h, w, 3)
    return out


def read_text(path):
    """-> {name: float32 H x W x 3}, parsed the way ImageTexture::parse_file does (src/main.cu:58-90)"""
    out = {}
    with open(path) as f:
        lines = f.read().split("\n")
    i = 1
    for _ in range(int(lines[0])):
        name, w, h = lines[i], int(lines[i + 1]), int(lines[i + 2])
        out[name] = np.array(lines[i + 3].split(), np.float64).astype(np.float32).reshape(h, w, 3)
        i += 4
    return out


def parse_directory(directory, binary=True):
    images = [(n, read_image(os.path.join(directory, n))) for n in list_images(directory)]
    write_text(os.path.join(directory, OUTPUT_TEXT), images)
    if binary:
        write_binary(os.path.join(directory, OUTPUT_BINARY), images)
    return images


if __name__ == "__main__":
    done = parse_directory(sys.argv[1] if len(sys.argv) > 1 else ".")
    print(f"{len(done)} image(s): " + ", ".join(f"{n} {p.shape[1]}x{p.shape[0]}" for n, p in done))

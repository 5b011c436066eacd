"""Procedural MDL v10 (Half-Life studio model) writer: synthetic skinned models for the ref_cuda benches and tests.

Layout as in engine/studio.h:145-420 of the reference: studiohdr_t, bones, one sequence (one blend) with RLE-packed
animation values per bone and channel (mstudioanim_t / mstudioanimvalue_t, decoded by R_StudioCalcBones,
public/xash3d_mathlib.c:423), one sequence group, one 8-bit indexed skin with its palette, one skin family, one
bodypart with one submodel: vertices and normals bound to single bones, meshes as triangle-strip commands
(count, then {vertex, normal, s, t} per strip vertex; 0 ends the list).

The model is a tube wound around a chain of `nbones` bones: rings of `ring` vertices at every joint (~1500 vertices at
32 x 48), two meshes (even / odd segments), one 30-frame looping sequence that bends the chain.  Everything derives
from the seed: the same arguments give the same bytes on every machine.
"""
from __future__ import annotations

import math
import struct

import numpy as np


def _name(s: str, n: int) -> bytes:
    b = s.encode("ascii")[: n - 1]
    return b + b"\0" * (n - len(b))


def generate(nbones: int = 32, ring: int = 48, seglen: float = 4.0, radius: float = 9.0, nframes: int = 30, skin: int = 128,
             seed: int = 0xB200, texflags: int = 0, name: str = "synth.mdl", attachments=(), bone_names=None) -> bytes:
    """attachments: (bone, (x, y, z)) pairs - mstudioattachment_t records (R_StudioCalcAttachments);
    bone_names: names of the first bones instead of bone00, bone01, ... (R_StudioMergeBones matches bones by name)"""
    rng = np.random.default_rng(seed)
    # ---- skin: 8-bit indices + palette
    ys, xs = np.mgrid[0:skin, 0:skin]
    idx = ((np.sin(xs * 0.37) * np.cos(ys * 0.23) * 0.5 + 0.5) * 180 + ((xs // 8 + ys // 8) % 2) * 60 + rng.integers(0, 14, (skin, skin))).astype(np.int64)
    idx = np.clip(idx, 0, 254).astype(np.uint8)
    t = (np.arange(256) / 255.0)[:, None]
    pal = np.clip(np.array([40.0, 30.0, 90.0])[None, :] * (1 - t) + np.array([250.0, 220.0, 140.0])[None, :] * t, 0, 255).astype(np.uint8)

    # ---- bones: a chain along local +z, root at the model origin
    bones = []
    for i in range(nbones):
        value = (0.0, 0.0, 0.0 if i == 0 else seglen, 0.0, 0.0, 0.0)
        scale = (0.0, 0.0, 0.0, 1.0 / 2048.0, 1.0 / 2048.0, 0.0)  # rotation about x and y is animated
        bones.append((bone_names[i] if bone_names and i < len(bone_names) else f"bone{i:02d}", i - 1, value, scale))

    # ---- geometry: vertex ring j of joint i lives in bone i's space
    verts, vbone = [], []
    for i in range(nbones):
        r = radius * (0.55 + 0.45 * math.cos(i * 0.41)) + 2.0
        for j in range(ring):
            a = 2 * math.pi * j / ring
            verts.append((r * math.cos(a), r * math.sin(a), 0.0))
            vbone.append(i)
    nverts = len(verts)
    # meshes: even and odd segments; every mesh owns its normals (one per ring vertex of the segments it draws)
    meshes = [[], []]
    for seg in range(nbones - 1):
        meshes[seg & 1].append(seg)
    norms, nbone, mesh_cmds, mesh_ntris, mesh_nnorms = [], [], [], [], []
    for segs in meshes:
        cmds, ntris, n0 = [], 0, len(norms)
        for seg in segs:
            base = len(norms)
            for ringi in (seg, seg + 1):
                for j in range(ring):
                    a = 2 * math.pi * j / ring
                    norms.append((math.cos(a), math.sin(a), 0.0))
                    nbone.append(ringi)
            cmds.append(2 * (ring + 1))  # one strip around the segment
            for j in range(ring + 1):
                jj = j % ring
                s = int(round(j * (skin - 1) / ring))
                for which, ringi in enumerate((seg, seg + 1)):
                    tt = int((ringi * 7) % (skin - 1))
                    cmds += [ringi * ring + jj, base + which * ring + jj, s, tt]
            ntris += 2 * ring
        cmds.append(0)
        mesh_cmds.append(cmds)
        mesh_ntris.append(ntris)
        mesh_nnorms.append(len(norms) - n0)

    # ---- animation: every bone sways about x and y, phase shifted along the chain
    anim_vals = np.zeros((nbones, 6, nframes), dtype=np.int16)
    for i in range(nbones):
        for f in range(nframes):
            ph = 2 * math.pi * f / nframes
            anim_vals[i, 3, f] = int(110 * math.sin(ph + i * 0.35))
            anim_vals[i, 4, f] = int(90 * math.cos(ph * 2 + i * 0.2))

    # ---- assemble
    blob = bytearray(b"\0" * 244)  # studiohdr_t
    def align():
        while len(blob) % 4:
            blob.append(0)

    boneindex = len(blob)
    for (bname, parent, value, scale) in bones:
        blob += _name(bname, 32) + struct.pack("<ii6i6f6f", parent, 0, *([-1] * 6), *value, *scale)
    seqindex = len(blob)
    seq_at = len(blob)
    blob += b"\0" * 176  # mstudioseqdesc_t, filled below
    seqgroupindex = len(blob)
    blob += _name("default", 32) + _name("", 64) + struct.pack("<ii", 0, 0)
    align()
    # animation data: nbones mstudioanim_t, then the value streams
    animindex = len(blob)
    offs = np.zeros((nbones, 6), dtype=np.uint16)
    streams = bytearray()
    base_streams = nbones * 12
    for i in range(nbones):
        for c in range(6):
            if not anim_vals[i, c].any():
                continue
            pos = base_streams + len(streams)
            offs[i, c] = pos - i * 12  # relative to this bone's mstudioanim_t
            streams += struct.pack("<BB", nframes, nframes) + anim_vals[i, c].astype("<i2").tobytes()
    for i in range(nbones):
        blob += struct.pack("<6H", *[int(x) for x in offs[i]])
    blob += streams
    align()
    # sequence
    seq = _name("sway", 32) + struct.pack("<f", 10.0) + struct.pack("<6i", 1, 0, 0, 0, 0, nframes)  # fps, flags=LOOPING, activity, actweight, numevents, eventindex, numframes
    seq += struct.pack("<2i", 0, 0)  # weightlistindex, iklockindex
    seq += struct.pack("<2i3f2i", 0, 0, 0.0, 0.0, 0.0, 0, 0)  # motiontype, motionbone, linearmovement, autolayerindex, keyvalueindex
    seq += struct.pack("<6f", -32, -32, 0, 32, 32, nbones * seglen)  # bbmin, bbmax
    seq += struct.pack("<2i", 1, animindex)  # numblends, animindex
    seq += struct.pack("<2i2f2f", 0, 0, 0.0, 0.0, 0.0, 0.0)  # blendtype, blendstart, blendend
    seq += struct.pack("<4B", 0, 0, 0, 0)  # groupsize[2], numautolayers, numiklocks
    seq += struct.pack("<3i", 0, 0, 0)  # seqgroup, entrynode, exitnode
    seq += struct.pack("<4B", 0, 0, 0, 0)  # nodeflags, cycleposeindex, fadeintime, fadeouttime
    seq += struct.pack("<i", 0)  # animdescindex
    assert len(seq) == 176, len(seq)
    blob[seq_at : seq_at + 176] = seq

    # bodypart + model + meshes
    bodypartindex = len(blob)
    blob += b"\0" * 76
    modelindex = len(blob)
    blob += b"\0" * 112
    vertinfoindex = len(blob)
    blob += bytes(vbone)
    align()
    norminfoindex = len(blob)
    blob += bytes(nbone)
    align()
    vertindex = len(blob)
    blob += np.array(verts, dtype="<f4").tobytes()
    normindex = len(blob)
    blob += np.array(norms, dtype="<f4").tobytes()
    meshindex = len(blob)
    blob += b"\0" * (20 * len(meshes))
    tri_at = []
    for cmds in mesh_cmds:
        tri_at.append(len(blob))
        blob += struct.pack(f"<{len(cmds)}h", *cmds)
        align()
    for k in range(len(meshes)):
        blob[meshindex + 20 * k : meshindex + 20 * k + 20] = struct.pack("<5i", mesh_ntris[k], tri_at[k], 0, mesh_nnorms[k], 0)
    blob[bodypartindex : bodypartindex + 76] = _name("body", 64) + struct.pack("<3i", 1, 1, modelindex)
    blob[modelindex : modelindex + 112] = _name("tube", 64) + struct.pack("<if10i", 0, 0.0, len(meshes), meshindex, nverts, vertinfoindex, vertindex,
                                                                           len(norms), norminfoindex, normindex, 0, 0)
    # texture header, skin table, texture data
    textureindex = len(blob)
    blob += b"\0" * 80
    skinindex = len(blob)
    blob += struct.pack("<h", 0)
    align()
    texturedataindex = len(blob)
    blob += idx.tobytes() + pal.tobytes()
    align()
    blob[textureindex : textureindex + 80] = _name("skin.bmp", 64) + struct.pack("<Iiii", texflags, skin, skin, texturedataindex)

    attachindex = len(blob)
    for k, (abone, aorg) in enumerate(attachments):
        blob += _name("attach%d" % k, 32) + struct.pack("<2i", 0, int(abone)) + struct.pack("<3f", *[float(v) for v in aorg]) + struct.pack("<9f", *([0.0] * 9))
    align()

    hdr = struct.pack("<ii", (ord("T") << 24) | (ord("S") << 16) | (ord("D") << 8) | ord("I"), 10) + _name(name, 64) + struct.pack("<i", len(blob))
    hdr += struct.pack("<3f", 0, 0, 0)  # eyeposition
    hdr += struct.pack("<6f", 0, 0, 0, 0, 0, 0)  # min, max
    hdr += struct.pack("<6f", -32, -32, 0, 32, 32, nbones * seglen)  # bbmin, bbmax
    hdr += struct.pack("<i", 0)  # flags
    hdr += struct.pack("<2i", nbones, boneindex)
    hdr += struct.pack("<2i", 0, boneindex)  # bone controllers: none
    hdr += struct.pack("<2i", 0, boneindex)  # hitboxes: none
    hdr += struct.pack("<2i", 1, seqindex)
    hdr += struct.pack("<2i", 1, seqgroupindex)
    hdr += struct.pack("<3i", 1, textureindex, texturedataindex)
    hdr += struct.pack("<3i", 1, 1, skinindex)
    hdr += struct.pack("<2i", 1, bodypartindex)
    hdr += struct.pack("<2i", len(attachments), attachindex if attachments else boneindex)
    hdr += struct.pack("<4i", 0, 0, 0, 0)  # studiohdr2index, unused x3
    hdr += struct.pack("<2i", 0, 0)  # transitions
    assert len(hdr) == 244, len(hdr)
    blob[0:244] = hdr
    return bytes(blob)


if __name__ == "__main__":
    import sys

    data = generate()
    print(len(data), "bytes")
    if len(sys.argv) > 1:
        open(sys.argv[1], "wb").write(data)

"""Canonical digest of a flattened scene (rtw_scene_desc): every value a kernel can see, with the
indices resolved (a primitive's transform matrices, its material, that material's texture, the
Perlin table / texels behind it), so that two descriptions which number their transforms,
materials and textures differently but describe the same scene give the same digest.  Used to
compare the product's host scene builder with the scene the REFERENCE'S OWN host code builds
(oracle/ref_py.scene_desc), tests/test_ref_pin.py."""
import ctypes as C
import hashlib
import struct


def resolved_records(rtw, desc_ptr):
    d = C.cast(desc_ptr, C.POINTER(rtw.SceneDesc)).contents
    table_hash = {}

    def blob_hash(kind, idx, data):
        key = (kind, idx)
        if key not in table_hash:
            table_hash[key] = hashlib.sha256(data).digest()
        return table_hash[key]

    recs = []
    for i in range(d.n_prims):
        p = d.prims[i]
        r = [struct.pack("<ii", p.type, p.flip), bytes(p.v)]
        r.append(bytes(d.transforms[p.transform]) if p.transform >= 0 else b"identity")
        m = d.materials[p.material]
        r.append(struct.pack("<if", m.type, m.param))
        if m.texture >= 0:
            t = d.textures[m.texture]
            r.append(struct.pack("<i", t.type))
            if t.type == rtw.TEX_CONSTANT:
                r.append(bytes(t.color))
            elif t.type == rtw.TEX_NOISE:
                r.append(struct.pack("<f", t.scale))
                r.append(blob_hash("perlin", t.perlin, bytes(d.perlin[t.perlin])))
            elif t.type == rtw.TEX_IMAGE:
                im = d.images[t.image]
                r.append(struct.pack("<ii", im.width, im.height))
                r.append(blob_hash("image", t.image, C.string_at(im.rgba, im.width * im.height * 4)))
            elif t.type == rtw.TEX_CHECKER:
                for c in (t.odd, t.even):
                    ct = d.textures[c]
                    r.append(struct.pack("<i", ct.type) + bytes(ct.color))
        else:
            r.append(b"notexture")
        recs.append(b"|".join(r))
    tail = [struct.pack("<ii", d.n_lights, d.sky_light)]
    for i in range(d.n_lights):
        tail.append(bytes(d.lights[i]))
    if d.n_lights > 0 and d.has_light_rect:   # without lights the rectangle is never read
        tail.append(bytes(d.light_rect))
    tail.append(bytes(d.camera))
    return recs, b"|".join(tail)


def scene_digest(rtw, desc_ptr):
    recs, tail = resolved_records(rtw, desc_ptr)
    h = hashlib.sha256()
    for r in recs:
        h.update(hashlib.sha256(r).digest())
    h.update(tail)
    return h.hexdigest()

"""Mitsuba 3 scene XML -> scene dictionary (the subset the hot path needs).

Follows the reference's parser where the hot-path scenes exercise it (mitsuba3/src/core/xml.cpp): ``<default name value>``
with ``$name`` substitution in every attribute (xml.cpp: parameter substitution happens before interpretation), plugin
elements (``scene integrator sensor sampler film rfilter bsdf texture shape emitter``) with ``type`` / ``id`` / ``name``,
property elements (``float integer boolean string rgb spectrum point vector``), ``<ref id name>``, ``<include filename>``,
``<path value>`` search directories, and ``<transform>`` with
``matrix translate rotate scale lookat`` composed left to right as ``T = op_n ... op_2 op_1`` (each new operation is applied
after the previous ones, xml.cpp transform handling).  The result has the schema of ``mi.load_dict`` / ``mi.cornell_box()``
(util.py:568-703) and goes straight into ``SceneDesc`` / ``load_dict``.

    d = load_file("data/bedroom/scene.xml", spp=16, resx=640, resy=360)
    scene = load_dict(d, base_path=os.path.dirname(path))
"""
from __future__ import annotations

import math
import os
import re
import xml.etree.ElementTree as ET

import numpy as np

from .scene import Transform4f

PLUGIN_TAGS = {"integrator", "sensor", "sampler", "film", "rfilter", "bsdf", "texture", "shape", "emitter", "medium", "phase", "volume"}
_VAR = re.compile(r"\$(\w+)")


class XMLError(ValueError):
    pass


def _floats(s: str) -> list[float]:
    return [float(x) for x in re.split(r"[\s,]+", s.strip()) if x]


def _transform(el, subst) -> np.ndarray:
    """<transform>: the child operations are composed in document order, each applied after the ones before it."""
    t = np.eye(4)
    for op in el:
        a = {k: subst(v) for k, v in op.attrib.items()}
        if op.tag == "matrix":
            v = _floats(a["value"])
            if len(v) == 9:
                m3 = np.asarray(v).reshape(3, 3); m = np.eye(4); m[:3, :3] = m3
            elif len(v) == 16:
                m = np.asarray(v).reshape(4, 4)
            else:
                raise XMLError("<matrix>: expected 9 or 16 values")
        elif op.tag == "translate":
            m = Transform4f().translate(_xyz(a, 0.0)).matrix
        elif op.tag == "scale":
            m = Transform4f().scale(_xyz(a, 1.0)).matrix
        elif op.tag == "rotate":
            axis = _xyz(a, 0.0)
            if not any(axis):
                raise XMLError("<rotate>: an axis (x / y / z or value) is required")
            m = Transform4f().rotate(axis, float(a["angle"])).matrix
        elif op.tag in ("lookat", "look_at"):
            m = Transform4f().look_at(_floats(a["origin"]), _floats(a["target"]), _floats(a.get("up", "0, 1, 0"))).matrix
        else:
            raise XMLError(f"<transform>: unsupported operation <{op.tag}>")
        t = m @ t
    return t


def _xyz(a: dict, default: float) -> list[float]:
    if "value" in a:
        v = _floats(a["value"])
        return v * 3 if len(v) == 1 else v
    return [float(a.get(k, default)) for k in ("x", "y", "z")]


class _Parser:
    def __init__(self, root, overrides):
        self.defaults = {}
        for el in root.iter("default"):
            self.defaults.setdefault(el.attrib["name"], el.attrib["value"])
        unknown = set(overrides) - set(self.defaults)
        if unknown:
            raise XMLError(f"unknown scene parameter(s): {sorted(unknown)} (declared: {sorted(self.defaults)})")
        self.defaults.update({k: str(v) for k, v in overrides.items()})
        self.id_tag = {el.attrib["id"]: el.tag for el in root.iter() if "id" in el.attrib and el.tag != "ref"}
        self.counter = 0

    def subst(self, s: str) -> str:
        def rep(m):
            if m.group(1) not in self.defaults:
                raise XMLError(f"undefined parameter ${m.group(1)}")
            return self.defaults[m.group(1)]
        return _VAR.sub(rep, s)

    def prop(self, el):
        a = {k: self.subst(v) for k, v in el.attrib.items()}
        t = el.tag
        if t == "float":
            return float(a["value"])
        if t == "integer":
            return int(a["value"])
        if t == "boolean":
            return a["value"].strip().lower() == "true"
        if t == "string":
            return a["value"]
        if t == "rgb":
            v = _floats(a["value"])
            return {"type": "rgb", "value": v * 3 if len(v) == 1 else v}
        if t == "spectrum":
            if "filename" in a:
                raise XMLError("<spectrum filename=...> is not supported")
            if ":" in a["value"]:
                raise XMLError("<spectrum value=\"wavelength:value, ...\">: wavelength lists need the CIE observer tables "
                               "(spectrum_list_to_srgb); give the colour as <rgb> instead")
            v = _floats(a["value"])
            if len(v) != 1:
                raise XMLError("<spectrum>: only uniform values are supported in the rgb variants")
            return {"type": "spectrum", "value": v[0]}
        if t in ("point", "vector"):
            return _xyz(a, 0.0)
        if t == "transform":
            return _transform(el, self.subst)
        raise XMLError(f"unsupported property element <{t}>")

    def child_key(self, el, parent_tag, used):
        a = el.attrib
        if "name" in a:
            return self.subst(a["name"])
        tag = self.id_tag.get(self.subst(a.get("id", "")), el.tag) if el.tag == "ref" else el.tag
        if parent_tag == "scene":
            if "id" in a and el.tag != "ref":
                return self.subst(a["id"])
            if tag in ("sensor", "integrator") and tag not in used:
                return tag
        elif tag in ("sampler", "film", "rfilter", "emitter", "bsdf", "texture") and tag not in used:
            return tag
        self.counter += 1
        return f"{tag}_{self.counter}"

    def plugin(self, el):
        a = {k: self.subst(v) for k, v in el.attrib.items()}
        if "type" not in a and el.tag != "scene":
            raise XMLError(f"<{el.tag}> needs a type attribute")
        d = {"type": a.get("type", "scene")}
        for ch in el:
            if ch.tag in ("default", "alias", "include", "path"):
                if ch.tag == "include":
                    raise XMLError("<include> needs a base directory: use load_file, or load_string(..., base_dir=...)")
                if ch.tag == "alias":
                    raise XMLError("<alias> is not supported")
                continue
            key = self.child_key(ch, el.tag, d)
            if key in d:
                raise XMLError(f"<{el.tag}>: duplicate property {key!r}")
            if ch.tag == "ref":
                d[key] = {"type": "ref", "id": self.subst(ch.attrib["id"])}
            elif ch.tag in PLUGIN_TAGS:
                d[key] = self.plugin(ch)
            else:
                d[key] = self.prop(ch)
        return d


def _expand_includes(root, base_dir, subst, depth=0):
    """<include filename="..."/> (xml.cpp: the included file's <scene> children are spliced in place; nested includes
    resolve against the included file's own directory)."""
    if depth > 16:
        raise XMLError("<include>: nesting too deep (cycle?)")
    out = []
    for ch in list(root):
        if ch.tag != "include":
            out.append(ch); continue
        if base_dir is None:
            raise XMLError("<include> needs a base directory: use load_file, or load_string(..., base_dir=...)")
        fname = subst(ch.attrib["filename"])
        path = fname if os.path.isabs(fname) else os.path.join(base_dir, fname)
        if not os.path.exists(path):
            raise XMLError(f"<include>: file {path!r} not found")
        inc = ET.parse(path).getroot()
        if inc.tag != "scene":
            raise XMLError(f"<include>: root element of {path!r} must be <scene>")
        _expand_includes(inc, os.path.dirname(path), subst, depth + 1)
        out.extend(list(inc))
    root[:] = out


def _resolve_filenames(d, base_dir, search):
    """<path value="..."/> search directories (xml.cpp / FileResolver): a relative ``filename`` that does not exist next to
    the scene file but does under a search path becomes absolute."""
    for k, v in d.items():
        if isinstance(v, dict):
            _resolve_filenames(v, base_dir, search)
        elif k == "filename" and isinstance(v, str) and not os.path.isabs(v) and not os.path.exists(os.path.join(base_dir, v)):
            for sp in search:
                if os.path.exists(os.path.join(sp, v)):
                    d[k] = os.path.abspath(os.path.join(sp, v)); break


def load_string(text: str, base_dir: str | None = None, **overrides) -> dict:
    """Parses scene XML into a scene dictionary; keyword arguments override ``<default>`` parameters (as ``mi.load_file`` does)."""
    root = ET.fromstring(text)
    if root.tag != "scene":
        raise XMLError("root element must be <scene>")
    # includes may use $parameters declared before them; fragments may declare further defaults
    pre = _Parser(root, {})
    pre.defaults.update({k: str(v) for k, v in overrides.items()})
    _expand_includes(root, base_dir, pre.subst)
    parser = _Parser(root, overrides)
    d = parser.plugin(root)
    d["type"] = "scene"
    if base_dir is not None:
        search = [parser.subst(el.attrib["value"]) for el in root.iter("path")]
        _resolve_filenames(d, base_dir, [p if os.path.isabs(p) else os.path.join(base_dir, p) for p in search])
    return d


def load_file(path: str, **overrides) -> dict:
    """``mi.load_file(path, **params)`` for the supported subset; relative filenames stay relative to ``dirname(path)``
    (pass it as ``base_path`` to ``SceneDesc`` / ``load_dict``)."""
    with open(path, "r", encoding="utf-8") as f:
        return load_string(f.read(), base_dir=os.path.dirname(os.path.abspath(path)), **overrides)

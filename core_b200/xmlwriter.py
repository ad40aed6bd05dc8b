"""Scene exporter to the reference's XML dialect — the role of ``xmlInterface_t``
(src/interface/xmlinterface.cc:30-330 of the reference): the same ``yafrayInterface_t`` method set, but every
call is written to an XML document that ``thebounty-xml`` (reference) or ``yb200-xml`` (tools/yb200_xml.cc) reads.

Any scene generator of ``core_b200.scenes`` can be pointed at an ``XmlWriter`` instead of a renderer.  Unlike the
reference writer, which prints 6 significant digits (xmlinterface.cc:112), numbers are written with round-trip
precision by default so that an XML-fed render is bit-identical to an API-fed one; pass ``digits=6`` for the
reference's formatting.
"""
import numpy as np


class XmlWriter:
    def __init__(self, path, digits=None):
        self._f = open(path, "w")
        self._digits = digits
        self._params = []          # current map: list of (name, xml-attribute string)
        self._eparams = []         # list of maps (shader-node lists)
        self._cur = self._params
        self._next_id = 2147483647  # scene_t::getNextFreeID counts down from INT_MAX (scene.cc:53,950-966)
        self._mat_names = {}
        self._cur_mat = None
        self._mesh_has_uv = False
        self._f.write('<?xml version="1.0"?>\n')

    # ---- number formatting
    def _n(self, v):
        v = float(v)
        if self._digits:
            return "%.*g" % (self._digits, v)
        return repr(v)

    # ---- scene
    def loadPlugins(self, path=None): pass
    def setVerbosityLevel(self, lvl): pass
    def setVerbosityMute(self): pass

    def startScene(self, type=0):
        self._f.write('<scene type="%s">\n' % ("triangle" if type == 0 else "universal"))
        return True

    def startGeometry(self): return True
    def endGeometry(self): return True

    def getNextFreeID(self):
        self._next_id -= 1
        return self._next_id + 1

    def startTriMesh(self, id, vertices, triangles, hasOrco, hasUV=False, type=0):
        self._cur_mat = None
        self._mesh_has_uv = bool(hasUV)
        self._n_uv = 0          # addUV returns the index of the value inside the mesh (scene.cc:648-662)
        self._n_vert = 0
        self._f.write('\n<mesh id="%d" vertices="%d" faces="%d" has_orco="%s" has_uv="%s" type="%d">\n' % (
            id, vertices, triangles, "true" if hasOrco else "false", "true" if hasUV else "false", type))
        return True

    def endTriMesh(self):
        self._f.write("</mesh>\n")
        return True

    def addVertex(self, x, y, z, ox=None, oy=None, oz=None):
        n = self._n
        if ox is None:
            self._f.write('\t<p x="%s" y="%s" z="%s"/>\n' % (n(x), n(y), n(z)))
        else:
            self._f.write('\t<p x="%s" y="%s" z="%s" ox="%s" oy="%s" oz="%s"/>\n' % (n(x), n(y), n(z), n(ox), n(oy), n(oz)))
        return 0

    def addNormal(self, nx, ny, nz):
        n = self._n
        self._f.write('\t<n x="%s" y="%s" z="%s"/>\n' % (n(nx), n(ny), n(nz)))

    def _set_material(self, mat):
        if mat != self._cur_mat:
            self._f.write('\t<set_material sval="%s"/>\n' % self._mat_names[mat])
            self._cur_mat = mat

    def addTriangle(self, a, b, c, *rest):
        if len(rest) == 1:
            self._set_material(rest[0])
            self._f.write('\t<f a="%d" b="%d" c="%d"/>\n' % (a, b, c))
        else:
            ua, ub, uc, mat = rest
            self._set_material(mat)
            self._f.write('\t<f a="%d" b="%d" c="%d" uv_a="%d" uv_b="%d" uv_c="%d"/>\n' % (a, b, c, ua, ub, uc))
        return True

    def addUV(self, u, v):
        self._f.write('\t<uv u="%s" v="%s"/>\n' % (self._n(u), self._n(v)))
        self._n_uv += 1
        return self._n_uv - 1

    def addVertices(self, xyz, normals=None):
        xyz = np.asarray(xyz, dtype=np.float32)
        for i in range(len(xyz)):
            self.addVertex(xyz[i, 0], xyz[i, 1], xyz[i, 2])
            if normals is not None:
                self.addNormal(*normals[i])
        return True

    def addTriangles(self, abc, mat):
        for t in np.asarray(abc, dtype=np.int32):
            self.addTriangle(int(t[0]), int(t[1]), int(t[2]), mat)
        return True

    def smoothMesh(self, id, angle):
        self._f.write('<smooth ID="%d" angle="%s"/>\n' % (id, self._n(angle)))
        return True

    def addInstance(self, baseObjectId, objToWorld):
        m = np.asarray(objToWorld, dtype=np.float32).reshape(4, 4)
        attrs = " ".join('m%d%d="%s"' % (i, j, self._n(m[i, j])) for i in range(4) for j in range(4))
        self._f.write('\n<instance base_object_id="%d">\n\t<transform %s/>\n</instance>\n' % (baseObjectId, attrs))
        return True

    # ---- parameters
    def paramsSetPoint(self, n, x, y, z): self._cur.append((n, 'x="%s" y="%s" z="%s"' % (self._n(x), self._n(y), self._n(z))))
    def paramsSetString(self, n, s): self._cur.append((n, 'sval="%s"' % s))
    def paramsSetBool(self, n, b): self._cur.append((n, 'bval="%s"' % ("true" if b else "false")))
    def paramsSetInt(self, n, i): self._cur.append((n, 'ival="%d"' % i))
    def paramsSetFloat(self, n, f): self._cur.append((n, 'fval="%s"' % self._n(f)))

    def paramsSetColor(self, n, r, g, b, a=1.0):
        self._cur.append((n, 'r="%s" g="%s" b="%s" a="%s"' % (self._n(r), self._n(g), self._n(b), self._n(a))))

    def paramsClearAll(self):
        self._params = []
        self._eparams = []
        self._cur = self._params

    def paramsStartList(self):
        if self._eparams:
            self._cur = self._eparams[-1]
        else:
            self.paramsPushList()

    def paramsPushList(self):
        self._eparams.append([])
        self._cur = self._eparams[-1]

    def paramsEndList(self):
        self._cur = self._params

    def _write_map(self, pm, indent="\t"):
        for k, v in pm:
            self._f.write("%s<%s %s/>\n" % (indent, k, v))

    def _element(self, kind, name):
        self._f.write('\n<%s name="%s">\n' % (kind, name))
        self._write_map(self._params)
        for sub in self._eparams:
            self._f.write("\t<list_element>\n")
            self._write_map(sub, "\t\t")
            self._f.write("\t</list_element>\n")
        self._f.write("</%s>\n" % kind)
        return name

    def createLight(self, name): return self._element("light", name)
    def createTexture(self, name): return self._element("texture", name)
    def createCamera(self, name): return self._element("camera", name)
    def createBackground(self, name): return self._element("background", name)
    def createIntegrator(self, name): return self._element("integrator", name)
    def createVolumeRegion(self, name): return self._element("volumeregion", name)

    def createMaterial(self, name):
        self._element("material", name)
        h = len(self._mat_names) + 1
        self._mat_names[h] = name
        return h

    # ---- render: writes the <render> map and closes the document (xmlinterface.cc:300-317)
    def render(self, width=None, height=None):
        self._f.write("\n<render>\n")
        self._write_map(self._params)
        self._f.write("</render>\n</scene>\n")
        self._f.close()
        return None

    def clearAll(self): pass
    def close(self):
        if not self._f.closed:
            self._f.close()

/*
 * t4_scene_compiler.cpp -- scene compiler from Tesseract's own input files to the PODs of include/t4c.h (SURVEY 8(f)
 * rank 2).  Host code only (no CUDA): built into libt4c_compiler.so.
 *
 *   URDF subset   links (<collision>: origin + box / sphere / cylinder / capsule / mesh with scale), joints (fixed,
 *                 revolute, continuous, prismatic: origin xyz / rpy, axis, limits)      tesseract_urdf/src/
 *   SRDF subset   <disable_collisions link1 link2>                                       tesseract_scene_graph/src/srdf_model.cpp
 *   meshes        binary / ASCII STL, ASCII / binary-little-endian PLY                   tesseract_geometry/mesh_parser.h (assimp there)
 *   mesh -> hull  own incremental convex hull; the hull VERTEX SET becomes the ConvexMesh
 *                 (makeConvexMesh / createConvexHull, tesseract_collision/core/common.h:181-264; Bullet's
 *                 btConvexHullComputer additionally snaps the vertices to an integer grid -- not reproduced)
 *   octomap .bt   binary tree stream -> occupied leaves at max depth                     bullet_utils.cpp:160-243, SURVEY A.9
 *   shape rules   one record per convex child: mesh -> triangle per face, octree -> box / sphere per leaf,
 *                 shape_id / subshape_id numbering                                       bullet_utils.cpp:59-365
 *   active links  everything below a moving joint                                        tesseract_environment/core/utils.h:55-76
 *
 * Output: a t4c_compiled object that owns the arrays behind a t4c_scene_desc and a t4c_fk_model_desc.
 */
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <string>
#include <vector>

#include "../../../include/t4c_compiler.h"

namespace
{
thread_local std::string g_cerr;
t4c_status cfail(t4c_status st, const std::string& msg)
{
  g_cerr = msg;
  return st;
}

// ---------------------------------------------------------------------------------------------------------------
// minimal XML reader: elements, attributes, nesting; comments, declarations and text are skipped
// ---------------------------------------------------------------------------------------------------------------
struct Xml
{
  std::string tag;
  std::map<std::string, std::string> attr;
  std::vector<std::unique_ptr<Xml>> kids;
  const Xml* child(const std::string& t) const
  {
    for (auto& k : kids)
      if (k->tag == t)
        return k.get();
    return nullptr;
  }
  std::vector<const Xml*> children(const std::string& t) const
  {
    std::vector<const Xml*> out;
    for (auto& k : kids)
      if (k->tag == t)
        out.push_back(k.get());
    return out;
  }
  std::string get(const std::string& a, const std::string& dflt = "") const
  {
    auto it = attr.find(a);
    return it == attr.end() ? dflt : it->second;
  }
};

struct XmlParser
{
  const std::string& s;
  size_t i = 0;
  explicit XmlParser(const std::string& src) : s(src) {}
  void skipWs()
  {
    while (i < s.size() && std::isspace((unsigned char)s[i]))
      ++i;
  }
  bool skipMisc()  // comments, <?...?>, <!...>, text
  {
    for (;;)
    {
      while (i < s.size() && s[i] != '<')
        ++i;
      if (i >= s.size())
        return false;
      if (s.compare(i, 4, "<!--") == 0)
      {
        const size_t e = s.find("-->", i);
        if (e == std::string::npos)
          return false;
        i = e + 3;
        continue;
      }
      if (s.compare(i, 2, "<?") == 0 || s.compare(i, 2, "<!") == 0)
      {
        const size_t e = s.find('>', i);
        if (e == std::string::npos)
          return false;
        i = e + 1;
        continue;
      }
      return true;
    }
  }
  std::string name()
  {
    const size_t a = i;
    while (i < s.size() && (std::isalnum((unsigned char)s[i]) || s[i] == '_' || s[i] == ':' || s[i] == '-' || s[i] == '.'))
      ++i;
    return s.substr(a, i - a);
  }
  std::unique_ptr<Xml> element()
  {
    if (!skipMisc() || s[i] != '<' || (i + 1 < s.size() && s[i + 1] == '/'))
      return nullptr;
    ++i;
    std::unique_ptr<Xml> e(new Xml());
    e->tag = name();
    for (;;)
    {
      skipWs();
      if (i >= s.size())
        return nullptr;
      if (s[i] == '/')
      {
        i += 2;  // "/>"
        return e;
      }
      if (s[i] == '>')
      {
        ++i;
        break;
      }
      const std::string key = name();
      skipWs();
      if (i >= s.size() || s[i] != '=')
        return nullptr;
      ++i;
      skipWs();
      const char q = s[i++];
      const size_t a = i;
      while (i < s.size() && s[i] != q)
        ++i;
      e->attr[key] = s.substr(a, i - a);
      ++i;
    }
    for (;;)
    {
      if (!skipMisc())
        return nullptr;
      if (s.compare(i, 2, "</") == 0)
      {
        const size_t c = s.find('>', i);
        i = c == std::string::npos ? s.size() : c + 1;
        return e;
      }
      std::unique_ptr<Xml> k = element();
      if (!k)
        return nullptr;
      e->kids.push_back(std::move(k));
    }
  }
};

bool readFile(const std::string& path, std::string& out)
{
  std::ifstream f(path, std::ios::binary);
  if (!f)
    return false;
  std::ostringstream ss;
  ss << f.rdbuf();
  out = ss.str();
  return true;
}

std::vector<double> numbers(const std::string& s)
{
  std::vector<double> v;
  std::istringstream is(s);
  double x;
  while (is >> x)
    v.push_back(x);
  return v;
}

// ---------------------------------------------------------------------------------------------------------------
// small linear algebra on 3x4 row-major poses
// ---------------------------------------------------------------------------------------------------------------
struct Pose
{
  double m[12];
  Pose()
  {
    static const double I[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
    std::memcpy(m, I, sizeof(m));
  }
  Pose operator*(const Pose& o) const
  {
    Pose r;
    for (int i = 0; i < 3; ++i)
    {
      for (int j = 0; j < 3; ++j)
        r.m[4 * i + j] = m[4 * i] * o.m[j] + m[4 * i + 1] * o.m[4 + j] + m[4 * i + 2] * o.m[8 + j];
      r.m[4 * i + 3] = m[4 * i] * o.m[3] + m[4 * i + 1] * o.m[7] + m[4 * i + 2] * o.m[11] + m[4 * i + 3];
    }
    return r;
  }
  bool identity() const
  {
    static const Pose I;
    return std::memcmp(m, I.m, sizeof(m)) == 0;
  }
};

// URDF origin: xyz + fixed-axis roll / pitch / yaw, R = Rz(yaw) Ry(pitch) Rx(roll)   (tesseract_urdf/src/origin.cpp)
Pose originOf(const Xml* o)
{
  Pose P;
  if (!o)
    return P;
  const std::vector<double> xyz = numbers(o->get("xyz", "0 0 0")), rpy = numbers(o->get("rpy", "0 0 0"));
  const double r = rpy.size() > 0 ? rpy[0] : 0, p = rpy.size() > 1 ? rpy[1] : 0, y = rpy.size() > 2 ? rpy[2] : 0;
  const double cr = std::cos(r), sr = std::sin(r), cp = std::cos(p), sp = std::sin(p), cy = std::cos(y), sy = std::sin(y);
  P.m[0] = cy * cp, P.m[1] = cy * sp * sr - sy * cr, P.m[2] = cy * sp * cr + sy * sr;
  P.m[4] = sy * cp, P.m[5] = sy * sp * sr + cy * cr, P.m[6] = sy * sp * cr - cy * sr;
  P.m[8] = -sp, P.m[9] = cp * sr, P.m[10] = cp * cr;
  for (int k = 0; k < 3; ++k)
    P.m[4 * k + 3] = xyz.size() > (size_t)k ? xyz[k] : 0.0;
  return P;
}

// ---------------------------------------------------------------------------------------------------------------
// meshes
// ---------------------------------------------------------------------------------------------------------------
struct TriMesh
{
  std::vector<double> v;   // xyz
  std::vector<int32_t> f;  // triples
};

void dedupe(const std::vector<float>& soup, TriMesh& out)  // soup: 9 floats per triangle
{
  std::map<std::array<float, 3>, int32_t> index;
  for (size_t t = 0; t < soup.size() / 3; ++t)
  {
    const std::array<float, 3> key = { soup[3 * t], soup[3 * t + 1], soup[3 * t + 2] };
    auto it = index.find(key);
    if (it == index.end())
    {
      it = index.emplace(key, (int32_t)(out.v.size() / 3)).first;
      out.v.push_back(key[0]), out.v.push_back(key[1]), out.v.push_back(key[2]);
    }
    out.f.push_back(it->second);
  }
}

bool loadStl(const std::string& data, TriMesh& out)
{
  std::vector<float> soup;
  const bool ascii = data.compare(0, 5, "solid") == 0 && data.find("facet", 0) != std::string::npos &&
                     data.find("facet", 0) < 512;
  if (ascii)
  {
    std::istringstream is(data);
    std::string tok;
    while (is >> tok)
      if (tok == "vertex")
      {
        float x, y, z;
        is >> x >> y >> z;
        soup.push_back(x), soup.push_back(y), soup.push_back(z);
      }
  }
  else
  {
    if (data.size() < 84)
      return false;
    uint32_t n;
    std::memcpy(&n, data.data() + 80, 4);
    if (data.size() < 84 + (size_t)n * 50)
      return false;
    soup.resize((size_t)n * 9);
    for (uint32_t t = 0; t < n; ++t)
      std::memcpy(&soup[(size_t)t * 9], data.data() + 84 + (size_t)t * 50 + 12, 36);
  }
  dedupe(soup, out);
  return !out.f.empty();
}

bool loadPly(const std::string& data, TriMesh& out)
{
  const size_t he = data.find("end_header");
  if (data.compare(0, 3, "ply") != 0 || he == std::string::npos)
    return false;
  std::istringstream hs(data.substr(0, he));
  std::string line;
  int nv = 0, nf = 0;
  bool binary = false, in_vertex = false;
  std::vector<std::string> vprops;
  while (std::getline(hs, line))
  {
    std::istringstream ls(line);
    std::string a, b, c;
    ls >> a >> b >> c;
    if (a == "format")
      binary = b != "ascii";
    else if (a == "element")
    {
      in_vertex = b == "vertex";
      if (b == "vertex")
        nv = std::atoi(c.c_str());
      else if (b == "face")
        nf = std::atoi(c.c_str());
    }
    else if (a == "property" && in_vertex)
      vprops.push_back(b);  // type of each vertex property, in order
  }
  size_t pos = data.find('\n', he);
  if (pos == std::string::npos)
    return false;
  ++pos;
  auto typeSize = [](const std::string& t) {
    if (t == "double" || t == "float64")
      return 8;
    if (t == "float" || t == "float32" || t == "int" || t == "uint" || t == "int32" || t == "uint32")
      return 4;
    if (t == "short" || t == "ushort" || t == "int16" || t == "uint16")
      return 2;
    return 1;
  };
  if (!binary)
  {
    std::istringstream bs(data.substr(pos));
    for (int i = 0; i < nv; ++i)
    {
      for (size_t p = 0; p < vprops.size(); ++p)
      {
        double x;
        bs >> x;
        if (p < 3)
          out.v.push_back((double)(float)x);
      }
    }
    for (int i = 0; i < nf; ++i)
    {
      int k;
      bs >> k;
      std::vector<int> idx((size_t)k);
      for (int j = 0; j < k; ++j)
        bs >> idx[j];
      for (int j = 1; j + 1 < k; ++j)
        out.f.push_back(idx[0]), out.f.push_back(idx[j]), out.f.push_back(idx[j + 1]);
    }
    return !out.f.empty();
  }
  for (int i = 0; i < nv; ++i)
    for (size_t p = 0; p < vprops.size(); ++p)
    {
      const int sz = typeSize(vprops[p]);
      if (pos + sz > data.size())
        return false;
      if (p < 3)
      {
        if (sz == 4)
        {
          float x;
          std::memcpy(&x, data.data() + pos, 4);
          out.v.push_back(x);
        }
        else
        {
          double x;
          std::memcpy(&x, data.data() + pos, 8);
          out.v.push_back(x);
        }
      }
      pos += sz;
    }
  for (int i = 0; i < nf; ++i)
  {
    if (pos >= data.size())
      return false;
    const int k = (unsigned char)data[pos++];
    std::vector<int32_t> idx((size_t)k);
    for (int j = 0; j < k; ++j)
    {
      std::memcpy(&idx[j], data.data() + pos, 4);
      pos += 4;
    }
    for (int j = 1; j + 1 < k; ++j)
      out.f.push_back(idx[0]), out.f.push_back(idx[j]), out.f.push_back(idx[j + 1]);
  }
  return !out.f.empty();
}

bool loadMesh(const std::string& path, TriMesh& out, std::string& err)
{
  std::string data;
  if (!readFile(path, data))
  {
    err = "cannot read mesh " + path;
    return false;
  }
  const std::string ext = path.size() > 4 ? path.substr(path.size() - 4) : "";
  const bool ok = (ext == ".ply" || ext == ".PLY") ? loadPly(data, out) : loadStl(data, out);
  if (!ok)
    err = "unsupported or empty mesh " + path + " (binary / ASCII STL and PLY are read)";
  return ok;
}

// ---------------------------------------------------------------------------------------------------------------
// incremental 3-D convex hull: returns the indices of the input points that are hull vertices (ascending)
// ---------------------------------------------------------------------------------------------------------------
struct V3
{
  double x, y, z;
};
inline V3 sub(const V3& a, const V3& b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline V3 cross(const V3& a, const V3& b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline double dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

std::vector<int32_t> hullVertices(const std::vector<double>& xyz)
{
  const int n = (int)(xyz.size() / 3);
  std::vector<V3> p((size_t)n);
  double scale = 0;
  for (int i = 0; i < n; ++i)
  {
    p[i] = { xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2] };
    scale = std::max(scale, std::max(std::fabs(p[i].x), std::max(std::fabs(p[i].y), std::fabs(p[i].z))));
  }
  std::vector<int32_t> all((size_t)n);
  for (int i = 0; i < n; ++i)
    all[i] = i;
  if (n < 4)
    return all;
  const double eps = 1e-12 * std::max(scale, 1e-30);
  // initial tetrahedron: extreme points
  int i0 = 0, i1 = -1, i2 = -1, i3 = -1;
  for (int i = 1; i < n; ++i)
    if (p[i].x < p[i0].x)
      i0 = i;
  double best = 0;
  for (int i = 0; i < n; ++i)
  {
    const V3 d = sub(p[i], p[i0]);
    if (dot(d, d) > best)
      best = dot(d, d), i1 = i;
  }
  if (i1 < 0)
    return { i0 };
  best = 0;
  for (int i = 0; i < n; ++i)
  {
    const V3 c = cross(sub(p[i1], p[i0]), sub(p[i], p[i0]));
    if (dot(c, c) > best)
      best = dot(c, c), i2 = i;
  }
  if (i2 < 0)
    return all;  // collinear input: nothing sensible to drop
  const V3 n012 = cross(sub(p[i1], p[i0]), sub(p[i2], p[i0]));
  best = 0;
  for (int i = 0; i < n; ++i)
  {
    const double d = std::fabs(dot(n012, sub(p[i], p[i0])));
    if (d > best)
      best = d, i3 = i;
  }
  if (i3 < 0 || best <= eps * std::sqrt(dot(n012, n012)))
    return all;  // coplanar input
  struct Face
  {
    int v[3];
    V3 n;
    double d;
    bool alive;
  };
  std::vector<Face> F;
  const V3 centre = { (p[i0].x + p[i1].x + p[i2].x + p[i3].x) / 4, (p[i0].y + p[i1].y + p[i2].y + p[i3].y) / 4,
                      (p[i0].z + p[i1].z + p[i2].z + p[i3].z) / 4 };
  auto addFace = [&](int a, int b, int c) {
    Face f;
    f.v[0] = a, f.v[1] = b, f.v[2] = c;
    f.n = cross(sub(p[b], p[a]), sub(p[c], p[a]));
    const double l = std::sqrt(dot(f.n, f.n));
    if (l > 0)
      f.n = { f.n.x / l, f.n.y / l, f.n.z / l };
    f.d = dot(f.n, p[a]);
    if (dot(f.n, centre) - f.d > 0)  // orient outwards
    {
      std::swap(f.v[1], f.v[2]);
      f.n = { -f.n.x, -f.n.y, -f.n.z };
      f.d = -f.d;
    }
    f.alive = true;
    F.push_back(f);
  };
  addFace(i0, i1, i2), addFace(i0, i1, i3), addFace(i0, i2, i3), addFace(i1, i2, i3);
  const double tol = 1e-10 * std::max(scale, 1e-30);
  for (int i = 0; i < n; ++i)
  {
    if (i == i0 || i == i1 || i == i2 || i == i3)
      continue;
    std::vector<int> vis;
    for (size_t f = 0; f < F.size(); ++f)
      if (F[f].alive && dot(F[f].n, p[i]) - F[f].d > tol)
        vis.push_back((int)f);
    if (vis.empty())
      continue;  // inside (or on) the current hull
    // horizon = directed edges of visible faces whose reverse is not an edge of a visible face
    std::set<std::pair<int, int>> edges;
    for (int f : vis)
      for (int e = 0; e < 3; ++e)
        edges.insert({ F[f].v[e], F[f].v[(e + 1) % 3] });
    std::vector<std::pair<int, int>> horizon;
    for (auto& e : edges)
      if (!edges.count({ e.second, e.first }))
        horizon.push_back(e);
    for (int f : vis)
      F[f].alive = false;
    for (auto& e : horizon)
      addFace(e.first, e.second, i);
  }
  std::set<int32_t> used;
  for (auto& f : F)
    if (f.alive)
      used.insert(f.v[0]), used.insert(f.v[1]), used.insert(f.v[2]);
  return std::vector<int32_t>(used.begin(), used.end());
}

// ---------------------------------------------------------------------------------------------------------------
// octomap .bt: occupied leaves in depth-first child order 0..7
// ---------------------------------------------------------------------------------------------------------------
bool loadBt(const std::string& data, std::vector<double>& leaves /* x y z size */, std::string& err)
{
  size_t pos = 0;
  double res = 0;
  long size = -1;
  for (;;)
  {
    const size_t e = data.find('\n', pos);
    if (e == std::string::npos)
    {
      err = "octomap header not terminated";
      return false;
    }
    const std::string line = data.substr(pos, e - pos);
    pos = e + 1;
    if (line.compare(0, 4, "res ") == 0)
      res = std::atof(line.c_str() + 4);
    else if (line.compare(0, 5, "size ") == 0)
      size = std::atol(line.c_str() + 5);
    else if (line == "data")
      break;
  }
  if (!(res > 0))
  {
    err = "octomap header has no resolution";
    return false;
  }
  // pre-order stream: every inner node is 2 bytes, 2 bits per child c at bits 2c (lo) / 2c + 1 (hi): lo=0 hi=1 occupied
  // leaf, lo=1 hi=0 free leaf, lo=1 hi=1 inner child (its own node follows, children in index order), 0 0 unknown.
  // Leaves come out in the order of octomap's leaf iterator: depth first, children 0..7.
  struct Walker
  {
    const std::string& data;
    size_t pos;
    double res;
    std::vector<double>& leaves;
    bool ok = true;
    void node(int depth, uint32_t kx, uint32_t ky, uint32_t kz)
    {
      if (pos + 2 > data.size())
      {
        ok = false;
        return;
      }
      const unsigned bits = (unsigned char)data[pos] | ((unsigned)(unsigned char)data[pos + 1] << 8);
      pos += 2;
      const int cd = depth + 1;
      const uint32_t half = 1u << (16 - cd);
      for (int c = 0; c < 8 && ok; ++c)
      {
        const unsigned lo = (bits >> (2 * c)) & 1u, hi = (bits >> (2 * c + 1)) & 1u;
        const uint32_t cx = kx + ((c & 1) ? half : 0), cy = ky + ((c & 2) ? half : 0), cz = kz + ((c & 4) ? half : 0);
        if (lo && hi)
          node(cd, cx, cy, cz);
        else if (!lo && hi)
        {
          const double edge = res * (double)half;
          leaves.push_back(((double)cx - 32768.0) * res + edge / 2);
          leaves.push_back(((double)cy - 32768.0) * res + edge / 2);
          leaves.push_back(((double)cz - 32768.0) * res + edge / 2);
          leaves.push_back(edge);
        }
      }
    }
  } w{ data, pos, res, leaves };
  w.node(0, 0, 0, 0);
  if (!w.ok)
  {
    err = "octomap stream ends early";
    return false;
  }
  long consumed = 0;
  (void)size;
  (void)consumed;
  return true;
}

}  // namespace

// ===================================================================================================================
// the compiled object
// ===================================================================================================================
struct t4c_compiled
{
  std::vector<std::string> link_names, joint_names;
  std::vector<t4c_link> links;
  std::vector<t4c_shape> shapes;
  std::vector<double> vertices;
  std::vector<int32_t> acm;
  std::vector<t4c_joint> joints;
  std::vector<double> limits;  // lower, upper per joint with a q column
  std::vector<int32_t> shapes_of_link;
  int n_q = 0, root = 0;
  double margin = 0;
  t4c_scene_desc sd;
  t4c_fk_model_desc fd;
  std::map<std::string, int> link_index;

  int addLink(const std::string& name)
  {
    auto it = link_index.find(name);
    if (it != link_index.end())
      return it->second;
    const int i = (int)links.size();
    link_index[name] = i;
    link_names.push_back(name);
    t4c_link l;
    std::memset(&l, 0, sizeof(l));
    l.enabled = 1;
    const Pose I;
    std::memcpy(l.pose, I.m, sizeof(l.pose));
    links.push_back(l);
    shapes_of_link.push_back(0);
    return i;
  }
  void shape(int type, int link, int shape_id, int sub, double p0, double p1, double p2, const Pose& P, int voff = 0,
             int vcount = 0)
  {
    t4c_shape s;
    std::memset(&s, 0, sizeof(s));
    s.type = type, s.link = link, s.shape_id = shape_id, s.subshape_id = sub;
    s.params[0] = p0, s.params[1] = p1, s.params[2] = p2;
    s.vertex_offset = voff, s.vertex_count = vcount;
    std::memcpy(s.pose, P.m, sizeof(s.pose));
    shapes.push_back(s);
  }
  void refresh()
  {
    for (size_t i = 0; i < links.size(); ++i)
      links[i].name = link_names[i].c_str();
    std::memset(&sd, 0, sizeof(sd));
    sd.n_links = (int)links.size(), sd.links = links.data();
    sd.n_shapes = (int)shapes.size(), sd.shapes = shapes.data();
    sd.n_vertices = (int)(vertices.size() / 3), sd.vertices = vertices.data();
    sd.n_acm_pairs = (int)(acm.size() / 2), sd.acm_pairs = acm.data();
    sd.default_margin = margin;
    std::memset(&fd, 0, sizeof(fd));
    fd.n_links = (int)links.size(), fd.n_joints = (int)joints.size(), fd.joints = joints.data();
    fd.n_q = n_q, fd.root_link = root;
  }
};

namespace
{
std::string resolve(const std::string& uri, const std::string& package_root)
{
  const std::string pre = "package://";
  if (uri.compare(0, pre.size(), pre) == 0)
    return package_root + "/" + uri.substr(pre.size());
  return uri;
}

// occupied leaves of an octomap -> one box / sphere per leaf, leaf-iterator order, child index = subshape_id
// (createShapePrimitive(Octree), bullet_utils.cpp:160-243: box half = size / 2, sphere inside r = size / 2, sphere
// outside r = sqrt(2) size / 2); `base` = pose of the octree in the link frame
void addOctomapLeaves(t4c_compiled& C, int li, int shape_id, const std::vector<double>& leaves, int sub_type, const Pose& base)
{
  for (size_t v = 0; v < leaves.size() / 4; ++v)
  {
    Pose L;
    L.m[3] = leaves[4 * v], L.m[7] = leaves[4 * v + 1], L.m[11] = leaves[4 * v + 2];
    const Pose P = base.identity() ? L : base * L;
    const double h = leaves[4 * v + 3] / 2;
    if (sub_type == 0)
      C.shape(T4C_BOX, li, shape_id, (int)v, h, h, h, P);
    else if (sub_type == 1)
      C.shape(T4C_SPHERE, li, shape_id, (int)v, h, 0, 0, P);
    else
      C.shape(T4C_SPHERE, li, shape_id, (int)v, std::sqrt(2 * (h * h)), 0, 0, P);
  }
}

// geometry of one <collision> -> shape records of link `li` (shape_id = index in the link's geometry list)
t4c_status addGeometry(t4c_compiled& C, int li, int shape_id, bool single, const Xml* geom, const Pose& P,
                       const std::string& package_root, bool convex)
{
  const int sub = single ? -1 : shape_id;  // compound child index becomes subshape_id (bullet_utils.h:422-423)
  if (const Xml* b = geom->child("box"))
  {
    const std::vector<double> s = numbers(b->get("size"));
    if (s.size() != 3)
      return cfail(T4C_ERR_INVALID, "box without a size");
    C.shape(T4C_BOX, li, shape_id, sub, s[0] / 2, s[1] / 2, s[2] / 2, P);
    return T4C_OK;
  }
  if (const Xml* s = geom->child("sphere"))
  {
    C.shape(T4C_SPHERE, li, shape_id, sub, std::atof(s->get("radius").c_str()), 0, 0, P);
    return T4C_OK;
  }
  if (const Xml* c = geom->child("cylinder"))
  {
    C.shape(T4C_CYLINDER, li, shape_id, sub, std::atof(c->get("radius").c_str()), std::atof(c->get("length").c_str()) / 2, 0, P);
    return T4C_OK;
  }
  if (const Xml* c = geom->child("capsule"))
  {
    C.shape(T4C_CAPSULE, li, shape_id, sub, std::atof(c->get("radius").c_str()), std::atof(c->get("length").c_str()) / 2, 0, P);
    return T4C_OK;
  }
  if (const Xml* c = geom->child("cone"))
  {
    C.shape(T4C_CONE, li, shape_id, sub, std::atof(c->get("radius").c_str()), std::atof(c->get("length").c_str()), 0, P);
    return T4C_OK;
  }
  if (const Xml* o = geom->child("octomap"))
  {
    // <octomap shape_type="box|sphere_inside|sphere_outside" [prune]><octree filename=".bt"/></octomap>
    // (tesseract_urdf/src/octomap.cpp:44-96, octree.cpp:33-59).  `prune` (merging of equal children after loading) is
    // not applied: a .bt file stores the tree as it was pruned when written.  <point_cloud> is not read.
    const std::string st = o->get("shape_type");
    const int sub_type = st == "box" ? 0 : (st == "sphere_inside" ? 1 : (st == "sphere_outside" ? 2 : -1));
    if (sub_type < 0)
      return cfail(T4C_ERR_INVALID, "octomap: shape_type must be 'box', 'sphere_inside' or 'sphere_outside'");
    const Xml* ot = o->child("octree");
    if (!ot)
      return cfail(o->child("point_cloud") ? T4C_ERR_UNSUPPORTED : T4C_ERR_INVALID,
                   "octomap: missing element 'octree' (point clouds are not read)");
    std::string data, err;
    const std::string path = resolve(ot->get("filename"), package_root);
    if (!readFile(path, data))
      return cfail(T4C_ERR_INVALID, "octree: cannot read " + path);
    std::vector<double> leaves;
    if (!loadBt(data, leaves, err))
      return cfail(T4C_ERR_INVALID, err);
    addOctomapLeaves(C, li, shape_id, leaves, sub_type, P);
    return T4C_OK;
  }
  const Xml* m = geom->child("mesh");
  const Xml* cm = geom->child("convex_mesh");
  if (m || cm)
  {
    const Xml* e = m ? m : cm;
    TriMesh mesh;
    std::string err;
    if (!loadMesh(resolve(e->get("filename"), package_root), mesh, err))
      return cfail(T4C_ERR_INVALID, err);
    std::vector<double> sc = numbers(e->get("scale", "1 1 1"));
    if (sc.size() != 3)
      sc = { 1, 1, 1 };
    for (size_t i = 0; i < mesh.v.size(); ++i)
      mesh.v[i] *= sc[i % 3];
    const int off = (int)(C.vertices.size() / 3);
    if (cm || convex)
    {
      const std::vector<int32_t> hv = hullVertices(mesh.v);
      for (int32_t i : hv)
        for (int k = 0; k < 3; ++k)
          C.vertices.push_back(mesh.v[3 * (size_t)i + k]);
      C.shape(T4C_HULL, li, shape_id, sub, 0, 0, 0, P, off, (int)hv.size());
      return T4C_OK;
    }
    // one triangle shape per face (bullet_utils.cpp:111-132): the triangle's own three vertices
    for (size_t f = 0; f < mesh.f.size() / 3; ++f)
    {
      for (int c = 0; c < 3; ++c)
        for (int k = 0; k < 3; ++k)
          C.vertices.push_back(mesh.v[3 * (size_t)mesh.f[3 * f + c] + k]);
      C.shape(T4C_TRIANGLE, li, shape_id, (int)f, 0, 0, 0, P, off + 3 * (int)f, 3);
    }
    return T4C_OK;
  }
  return cfail(T4C_ERR_UNSUPPORTED, "unsupported collision geometry in link " + C.link_names[li]);
}
}  // namespace

extern "C" {
const char* t4c_compiler_last_error(void) { return g_cerr.c_str(); }

t4c_status t4c_compile_urdf(const char* urdf_path, const char* srdf_path, const char* package_root,
                            int32_t meshes_as_convex_hulls, double default_margin, t4c_compiled** out)
{
  if (!urdf_path || !out)
    return cfail(T4C_ERR_INVALID, "null argument");
  std::string text;
  if (!readFile(urdf_path, text))
    return cfail(T4C_ERR_INVALID, std::string("cannot read ") + urdf_path);
  XmlParser xp(text);
  std::unique_ptr<Xml> robot = xp.element();
  if (!robot || robot->tag != "robot")
    return cfail(T4C_ERR_INVALID, "not a URDF: no <robot> element");
  std::unique_ptr<t4c_compiled> C(new t4c_compiled());
  C->margin = default_margin;
  const std::string root_dir = package_root ? package_root : "";
  for (const Xml* l : robot->children("link"))
  {
    const int li = C->addLink(l->get("name"));
    const std::vector<const Xml*> cols = l->children("collision");
    // a single shape with an identity pose is used directly, anything else becomes a compound (bullet_utils.cpp:334-360)
    const bool single = cols.size() == 1 && originOf(cols[0]->child("origin")).identity();
    int shape_id = 0;
    for (const Xml* c : cols)
    {
      const Xml* g = c->child("geometry");
      if (!g)
        return cfail(T4C_ERR_INVALID, "collision without geometry in link " + l->get("name"));
      const t4c_status st = addGeometry(*C, li, shape_id, single, g, originOf(c->child("origin")), root_dir,
                                        meshes_as_convex_hulls != 0);
      if (st != T4C_OK)
        return st;
      ++shape_id;
    }
  }
  // joints (parent before child order is established below)
  struct J
  {
    std::string name;
    t4c_joint j;
    double lo, hi;
  };
  std::vector<J> js;
  for (const Xml* jx : robot->children("joint"))
  {
    J e;
    std::memset(&e.j, 0, sizeof(e.j));
    e.name = jx->get("name");
    const std::string type = jx->get("type");
    e.j.type = type == "fixed" ? T4C_JOINT_FIXED : (type == "prismatic" ? T4C_JOINT_PRISMATIC : T4C_JOINT_REVOLUTE);
    if (type != "fixed" && type != "prismatic" && type != "revolute" && type != "continuous")
      return cfail(T4C_ERR_UNSUPPORTED, "joint type '" + type + "' of " + e.name);
    const Xml* p = jx->child("parent");
    const Xml* c = jx->child("child");
    if (!p || !c)
      return cfail(T4C_ERR_INVALID, "joint without parent / child: " + e.name);
    e.j.parent_link = C->addLink(p->get("link"));
    e.j.child_link = C->addLink(c->get("link"));
    const Pose O = originOf(jx->child("origin"));
    std::memcpy(e.j.origin, O.m, sizeof(e.j.origin));
    std::vector<double> ax = numbers(jx->child("axis") ? jx->child("axis")->get("xyz", "1 0 0") : "1 0 0");
    if (ax.size() != 3)
      ax = { 1, 0, 0 };
    for (int k = 0; k < 3; ++k)
      e.j.axis[k] = ax[k];
    e.lo = e.hi = 0;
    if (const Xml* lim = jx->child("limit"))
    {
      e.lo = std::atof(lim->get("lower", "0").c_str());
      e.hi = std::atof(lim->get("upper", "0").c_str());
    }
    if (type == "continuous")
      e.lo = -M_PI, e.hi = M_PI;
    e.j.q_index = -1;
    js.push_back(e);
  }
  // q columns in URDF joint order (the order Tesseract's kinematic groups and state solvers use for a serial chain)
  for (auto& e : js)
    if (e.j.type != T4C_JOINT_FIXED)
    {
      e.j.q_index = C->n_q++;
      C->limits.push_back(e.lo), C->limits.push_back(e.hi);
    }
  // parent-before-child order
  std::set<int> children;
  for (auto& e : js)
    children.insert(e.j.child_link);
  std::set<int> placed;
  for (int i = 0; i < (int)C->links.size(); ++i)
    if (!children.count(i))
      placed.insert(i);
  C->root = placed.empty() ? 0 : *placed.begin();
  std::vector<bool> done(js.size(), false);
  for (size_t round = 0; round < js.size(); ++round)
  {
    bool progressed = false;
    for (size_t k = 0; k < js.size(); ++k)
      if (!done[k] && placed.count(js[k].j.parent_link))
      {
        C->joints.push_back(js[k].j);
        C->joint_names.push_back(js[k].name);
        placed.insert(js[k].j.child_link);
        done[k] = true;
        progressed = true;
      }
    if (!progressed)
      break;
  }
  if (C->joints.size() != js.size())
    return cfail(T4C_ERR_INVALID, "joint graph is not a forest");
  // active links: everything below a moving joint (core/utils.h:55-76)
  std::vector<bool> moving(C->links.size(), false);
  for (auto& j : C->joints)
    if (j.type != T4C_JOINT_FIXED || moving[j.parent_link])
      moving[j.child_link] = true;
  for (size_t i = 0; i < C->links.size(); ++i)
    C->links[i].active = moving[i] ? 1 : 0;
  // static links that hang on fixed joints keep a world pose: compose it from the root
  {
    std::vector<Pose> world(C->links.size());
    for (auto& j : C->joints)
    {
      Pose O;
      std::memcpy(O.m, j.origin, sizeof(O.m));
      world[j.child_link] = world[j.parent_link] * O;  // q = 0 pose: exact for fixed joints, start pose otherwise
    }
    for (size_t i = 0; i < C->links.size(); ++i)
      std::memcpy(C->links[i].pose, world[i].m, sizeof(world[i].m));
  }
  if (srdf_path && *srdf_path)
  {
    std::string st;
    if (!readFile(srdf_path, st))
      return cfail(T4C_ERR_INVALID, std::string("cannot read ") + srdf_path);
    XmlParser sp(st);
    std::unique_ptr<Xml> srdf = sp.element();
    if (!srdf)
      return cfail(T4C_ERR_INVALID, "SRDF does not parse");
    for (const Xml* d : srdf->children("disable_collisions"))
    {
      auto a = C->link_index.find(d->get("link1")), b = C->link_index.find(d->get("link2"));
      if (a != C->link_index.end() && b != C->link_index.end())
        C->acm.push_back(a->second), C->acm.push_back(b->second);
    }
  }
  C->refresh();
  *out = C.release();
  return T4C_OK;
}

t4c_status t4c_compiled_add_octomap(t4c_compiled* c, const char* link_name, const char* bt_path, int32_t sub_type,
                                    const double* pose12)
{
  if (!c || !link_name || !bt_path)
    return cfail(T4C_ERR_INVALID, "null argument");
  std::string data, err;
  if (!readFile(bt_path, data))
    return cfail(T4C_ERR_INVALID, std::string("cannot read ") + bt_path);
  std::vector<double> leaves;
  if (!loadBt(data, leaves, err))
    return cfail(T4C_ERR_INVALID, err);
  const int li = c->addLink(link_name);
  if (pose12)
    std::memcpy(c->links[li].pose, pose12, 12 * sizeof(double));
  const int shape_id = c->shapes_of_link[li]++;
  addOctomapLeaves(*c, li, shape_id, leaves, sub_type == 0 ? 0 : (sub_type == 1 ? 1 : 2), Pose());
  c->refresh();
  return T4C_OK;
}

t4c_status t4c_compiled_add_object(t4c_compiled* c, const char* link_name, int32_t n_shapes, const t4c_shape* shapes,
                                   int32_t n_vertices, const double* vertices, const double* pose12)
{
  if (!c || !link_name || n_shapes < 0 || (n_shapes > 0 && !shapes) || n_vertices < 0 || (n_vertices > 0 && !vertices))
    return cfail(T4C_ERR_INVALID, "bad argument");
  const int li = c->addLink(link_name);
  if (pose12)
    std::memcpy(c->links[li].pose, pose12, 12 * sizeof(double));
  const int off = (int)(c->vertices.size() / 3);
  c->vertices.insert(c->vertices.end(), vertices, vertices + 3 * (size_t)n_vertices);
  for (int i = 0; i < n_shapes; ++i)
  {
    t4c_shape s = shapes[i];
    s.link = li;
    if (s.type == T4C_HULL || s.type == T4C_TRIANGLE)
      s.vertex_offset += off;
    c->shapes.push_back(s);
  }
  c->shapes_of_link[li] += n_shapes;
  c->refresh();
  return T4C_OK;
}

const t4c_scene_desc* t4c_compiled_scene(const t4c_compiled* c) { return c ? &c->sd : nullptr; }
const t4c_fk_model_desc* t4c_compiled_fk(const t4c_compiled* c) { return c ? &c->fd : nullptr; }
int32_t t4c_compiled_link_index(const t4c_compiled* c, const char* name)
{
  if (!c || !name)
    return -1;
  auto it = c->link_index.find(name);
  return it == c->link_index.end() ? -1 : it->second;
}
const char* t4c_compiled_joint_name(const t4c_compiled* c, int32_t joint)
{
  return (c && joint >= 0 && joint < (int)c->joint_names.size()) ? c->joint_names[joint].c_str() : nullptr;
}
t4c_status t4c_compiled_limits(const t4c_compiled* c, double* lower, double* upper)
{
  if (!c || !lower || !upper)
    return cfail(T4C_ERR_INVALID, "null argument");
  for (int i = 0; i < c->n_q; ++i)
    lower[i] = c->limits[2 * i], upper[i] = c->limits[2 * i + 1];
  return T4C_OK;
}
void t4c_compiled_destroy(t4c_compiled* c) { delete c; }

t4c_status t4c_convex_hull_vertices(const double* xyz, int32_t n, int32_t* indices, int32_t* n_out)
{
  if (!xyz || !indices || !n_out || n < 0)
    return cfail(T4C_ERR_INVALID, "bad argument");
  const std::vector<int32_t> hv = hullVertices(std::vector<double>(xyz, xyz + 3 * (size_t)n));
  *n_out = (int32_t)hv.size();
  std::copy(hv.begin(), hv.end(), indices);
  return T4C_OK;
}

t4c_status t4c_read_octomap_bt(const char* bt_path, int64_t capacity, double* leaves, int64_t* n_leaves)
{
  if (!bt_path || !n_leaves)
    return cfail(T4C_ERR_INVALID, "null argument");
  std::string data, err;
  if (!readFile(bt_path, data))
    return cfail(T4C_ERR_INVALID, std::string("cannot read ") + bt_path);
  std::vector<double> lv;
  if (!loadBt(data, lv, err))
    return cfail(T4C_ERR_INVALID, err);
  *n_leaves = (int64_t)(lv.size() / 4);
  if (leaves)
    std::copy(lv.begin(), lv.begin() + 4 * (size_t)std::min<int64_t>(capacity, *n_leaves), leaves);
  return T4C_OK;
}
}  // extern "C"

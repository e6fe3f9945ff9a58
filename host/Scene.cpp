// Scene.cpp — CLRayTracer::Scene::load re-hosted: scene description file -> packed v0.6 scene buffer.
//
// Follows R/src/CLRayTracerEngine/Scene.cpp:231-444 (R = /root/reference/OpenCLRayTracer):
//   * the file is a list of `KEY value...` lines; keys LIGHT (x y z energy), SPHERE (x y z r), MESH (path to a Wavefront OBJ);
//     lines are grouped per key (std::map order: LIGHT < MESH < SPHERE), other keys are ignored (:236-247);
//   * every MESH is loaded, its materials are de-duplicated against the scene's list with an epsilon compare and the
//     shapes' material ids are rewritten (:129-167); a shape becomes one sub-mesh whose material is that of its FIRST face (:411);
//   * buffer = SceneHeader | Light[] | Sphere[] | Material[] | per model: ModelHeader | per sub-mesh: MeshHeader |
//     float4 vertices | ushort indices (SceneBufferParser.h:59-153); sub-mesh blocks are not padded (Q14), indices are
//     truncated to 16 bits (:429-430, Q13); boxes start from (FLT_MAX.., FLT_MIN..) (:383-384,397-398, Q4).
//   * quirk kept: the running size adds the material table once per recognised key that follows the MESH key in map
//     order (:313), so a file with SPHERE lines over-allocates by one material table; the tail stays zero.
// The OBJ / MTL reader restates what the reference's vendored loader (tiny_obj_loader 0.9.17,
// R/Include/3rdParty/tiny_obj_loader.h) does with a file, because the packed bytes depend on it:
//   * numbers: its own decimal parser (:266-368) — integer digits accumulate exactly, fraction digits add
//     digit * pow(10, -k), the value is sign * ldexp(mantissa * pow(5, e), e), then narrowed to float;
//   * faces are fan-triangulated (:535-556); a shape is closed at every `g`, `o` and `usemtl` (:973-1049) and kept only if
//     it has faces; inside a shape vertices are de-duplicated on the (v, vn, vt) index triple (:454-491);
//   * `usemtl` of an unknown name gives material id -1, which the packer's map lookup turns into 0 (Scene.cpp:162-164).
// Windows path separators in MESH lines are accepted (the reference's scene files use them); a relative model path is
// tried against the working directory, the scene file's directory and its Release/ sub-directory (the reference's
// .bat files run from Bin/Release).
#include <clrt/CLRayTracer.hpp>
#include <clrt_host.h>

#include <cstdio>
#include <cstring>
#include <dirent.h>
#include <strings.h>
#include <cstdlib>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

using CLRayTracer::Common::Errata;
using CLRayTracer::Common::Result;

// ---- OBJ / MTL ---------------------------------------------------------------------------------------------------
struct ObjMaterial {
  std::string name;
  float ambient[3], diffuse[3], specular[3], transmittance[3], emission[3];
  float shininess, ior, dissolve;
  int illum;
  ObjMaterial() { reset(); }
  void reset() {
    name.clear();
    for (int i = 0; i < 3; i++) ambient[i] = diffuse[i] = specular[i] = transmittance[i] = emission[i] = 0.0f;
    illum = 0; dissolve = 1.0f; shininess = 1.0f; ior = 1.0f;
  }
};
struct ObjShape {
  std::vector<float> positions;          // xyz per de-duplicated vertex
  std::vector<unsigned int> indices;     // 3 per triangle
  std::vector<int> material_ids;         // per triangle
};
struct VertexKey {
  int v, vn, vt;
  bool operator<(const VertexKey& o) const { return v != o.v ? v < o.v : vn != o.vn ? vn < o.vn : vt < o.vt; }
};

inline bool is_space(char c) { return c == ' ' || c == '\t'; }
inline bool is_newline(char c) { return c == '\r' || c == '\n' || c == '\0'; }

// the loader's decimal grammar: [sign] digits [. digits] [(e|E) [sign] digits]; anything else leaves 0
double parse_decimal(const char* s, const char* end) {
  if (s >= end) return 0.0;
  double mantissa = 0.0;
  int exponent = 0;
  char sign = '+', exp_sign = '+';
  const char* c = s;
  if (*c == '+' || *c == '-') { sign = *c; c++; }
  else if (!isdigit((unsigned char)*c)) return 0.0;
  int read = 0;
  while (c != end && isdigit((unsigned char)*c)) { mantissa *= 10; mantissa += (int)(*c - '0'); c++; read++; }
  if (read == 0) return 0.0;
  bool more = c != end;
  if (more && *c == '.') {
    c++; read = 1;
    while (c != end && isdigit((unsigned char)*c)) { mantissa += (int)(*c - '0') * pow(10.0, -read); read++; c++; }
    more = c != end;
  } else if (more && !(*c == 'e' || *c == 'E')) {
    more = false;
  }
  if (more && (*c == 'e' || *c == 'E')) {
    c++;
    if (c != end && (*c == '+' || *c == '-')) { exp_sign = *c; c++; }
    else if (!(c != end && isdigit((unsigned char)*c))) return 0.0;
    read = 0;
    while (c != end && isdigit((unsigned char)*c)) { exponent *= 10; exponent += (int)(*c - '0'); c++; read++; }
    exponent *= (exp_sign == '+' ? 1 : -1);
    if (read == 0) return 0.0;
  }
  return (sign == '+' ? 1 : -1) * ldexp(mantissa * pow(5.0, exponent), exponent);
}
float parse_float(const char*& tok) {
  tok += strspn(tok, " \t");
  const char* end = tok + strcspn(tok, " \t\r");
  const float f = (float)parse_decimal(tok, end);
  tok = end;
  return f;
}
void parse_float3(float* out, const char*& tok) { out[0] = parse_float(tok); out[1] = parse_float(tok); out[2] = parse_float(tok); }
int parse_int(const char*& tok) {
  tok += strspn(tok, " \t");
  const int i = atoi(tok);
  tok += strcspn(tok, " \t\r");
  return i;
}
std::string first_word(const char* tok) {                 // sscanf("%s")
  tok += strspn(tok, " \t\r\n\v\f");
  return std::string(tok, strcspn(tok, " \t\r\n\v\f"));
}
inline int fix_index(int idx, int n) { return idx > 0 ? idx - 1 : idx == 0 ? 0 : n + idx; }
VertexKey parse_triple(const char*& tok, int vsize, int vnsize, int vtsize) {      // i, i/j, i//k, i/j/k
  VertexKey k; k.v = k.vn = k.vt = -1;
  k.v = fix_index(atoi(tok), vsize);
  tok += strcspn(tok, "/ \t\r");
  if (tok[0] != '/') return k;
  tok++;
  if (tok[0] == '/') { tok++; k.vn = fix_index(atoi(tok), vnsize); tok += strcspn(tok, "/ \t\r"); return k; }
  k.vt = fix_index(atoi(tok), vtsize);
  tok += strcspn(tok, "/ \t\r");
  if (tok[0] != '/') return k;
  tok++;
  k.vn = fix_index(atoi(tok), vnsize);
  tok += strcspn(tok, "/ \t\r");
  return k;
}

bool read_line(std::istream& in, std::string& line) {     // getline into an 8192-byte buffer, CR/LF trimmed
  if (in.peek() == -1) return false;
  std::getline(in, line);
  if (!line.empty() && line[line.size() - 1] == '\n') line.erase(line.size() - 1);
  if (!line.empty() && line[line.size() - 1] == '\r') line.erase(line.size() - 1);
  return true;
}

void load_mtl(std::istream& in, std::map<std::string, int>& by_name, std::vector<ObjMaterial>& materials) {
  ObjMaterial m;
  std::string line;
  while (read_line(in, line)) {
    if (line.empty()) continue;
    const char* tok = line.c_str();
    tok += strspn(tok, " \t");
    if (tok[0] == '\0' || tok[0] == '#') continue;
    if (0 == strncmp(tok, "newmtl", 6) && is_space(tok[6])) {
      if (!m.name.empty()) { by_name.insert(std::make_pair(m.name, (int)materials.size())); materials.push_back(m); }
      m.reset();
      m.name = first_word(tok + 7);
      continue;
    }
    if (tok[0] == 'K' && tok[1] == 'a' && is_space(tok[2])) { tok += 2; parse_float3(m.ambient, tok); continue; }
    if (tok[0] == 'K' && tok[1] == 'd' && is_space(tok[2])) { tok += 2; parse_float3(m.diffuse, tok); continue; }
    if (tok[0] == 'K' && tok[1] == 's' && is_space(tok[2])) { tok += 2; parse_float3(m.specular, tok); continue; }
    if (tok[0] == 'K' && tok[1] == 't' && is_space(tok[2])) { tok += 2; parse_float3(m.transmittance, tok); continue; }
    if (tok[0] == 'N' && tok[1] == 'i' && is_space(tok[2])) { tok += 2; m.ior = parse_float(tok); continue; }
    if (tok[0] == 'K' && tok[1] == 'e' && is_space(tok[2])) { tok += 2; parse_float3(m.emission, tok); continue; }
    if (tok[0] == 'N' && tok[1] == 's' && is_space(tok[2])) { tok += 2; m.shininess = parse_float(tok); continue; }
    if (0 == strncmp(tok, "illum", 5) && is_space(tok[5])) { tok += 6; m.illum = parse_int(tok); continue; }
    if (tok[0] == 'd' && is_space(tok[1])) { tok += 1; m.dissolve = parse_float(tok); continue; }
    if (tok[0] == 'T' && tok[1] == 'r' && is_space(tok[2])) { tok += 2; m.dissolve = 1.0f - parse_float(tok); continue; }
    // texture maps and unknown parameters do not reach the packed scene
  }
  by_name.insert(std::make_pair(m.name, (int)materials.size()));        // the last (or the default, unnamed) material
  materials.push_back(m);
}

void close_shape(ObjShape& shape, const std::vector<float>& v, const std::vector<std::vector<VertexKey> >& faces, int material,
                 std::vector<ObjShape>& shapes) {
  if (faces.empty()) { shape = ObjShape(); return; }
  std::map<VertexKey, unsigned int> cache;              // per shape: the loader passes its cache by value
  for (size_t f = 0; f < faces.size(); f++) {
    const std::vector<VertexKey>& face = faces[f];
    VertexKey i0 = face[0], i1, i2 = face[1];
    i1.v = i1.vn = i1.vt = -1;
    for (size_t k = 2; k < face.size(); k++) {
      i1 = i2; i2 = face[k];
      const VertexKey tri[3] = {i0, i1, i2};
      for (int c = 0; c < 3; c++) {
        std::map<VertexKey, unsigned int>::const_iterator it = cache.find(tri[c]);
        unsigned int idx;
        if (it != cache.end()) idx = it->second;
        else {
          const size_t b = 3 * (size_t)tri[c].v;
          shape.positions.push_back(v.at(b)); shape.positions.push_back(v.at(b + 1)); shape.positions.push_back(v.at(b + 2));
          idx = (unsigned int)(shape.positions.size() / 3 - 1);
          cache[tri[c]] = idx;
        }
        shape.indices.push_back(idx);
      }
      shape.material_ids.push_back(material);
    }
  }
  if (!shape.indices.empty() || true) shapes.push_back(shape);      // kept whenever the face group was not empty
  shape = ObjShape();
}

std::string ci_resolve(const std::string& p);

bool load_obj(const std::string& file, const std::string& mtl_base, std::vector<ObjShape>& shapes, std::vector<ObjMaterial>& materials,
              std::string& err) {
  std::ifstream in(file.c_str());
  if (!in) { err = "Cannot open file [" + file + "]"; return false; }
  std::vector<float> v; int vn = 0, vt = 0;
  std::vector<std::vector<VertexKey> > faces;
  std::map<std::string, int> by_name;
  int material = -1;
  ObjShape shape;
  std::string line;
  while (read_line(in, line)) {
    if (line.empty()) continue;
    const char* tok = line.c_str();
    tok += strspn(tok, " \t");
    if (tok[0] == '\0' || tok[0] == '#') continue;
    if (tok[0] == 'v' && is_space(tok[1])) { tok += 2; float p[3]; parse_float3(p, tok); v.push_back(p[0]); v.push_back(p[1]); v.push_back(p[2]); continue; }
    if (tok[0] == 'v' && tok[1] == 'n' && is_space(tok[2])) { vn++; continue; }
    if (tok[0] == 'v' && tok[1] == 't' && is_space(tok[2])) { vt++; continue; }
    if (tok[0] == 'f' && is_space(tok[1])) {
      tok += 2; tok += strspn(tok, " \t");
      std::vector<VertexKey> face;
      while (!is_newline(tok[0])) { face.push_back(parse_triple(tok, (int)(v.size() / 3), vn, vt)); tok += strspn(tok, " \t\r"); }
      if (face.size() < 2) { err = "face with fewer than two vertices in [" + file + "]"; return false; }
      faces.push_back(face);
      continue;
    }
    if (0 == strncmp(tok, "usemtl", 6) && is_space(tok[6])) {
      const std::string name = first_word(tok + 7);
      close_shape(shape, v, faces, material, shapes); faces.clear();
      std::map<std::string, int>::const_iterator it = by_name.find(name);
      material = it != by_name.end() ? it->second : -1;
      continue;
    }
    if (0 == strncmp(tok, "mtllib", 6) && is_space(tok[6])) {
      const std::string path = ci_resolve(mtl_base + first_word(tok + 7));
      std::ifstream m(path.c_str());
      load_mtl(m, by_name, materials);                  // a missing file still yields the default material
      continue;
    }
    if ((tok[0] == 'g' || tok[0] == 'o') && is_space(tok[1])) { close_shape(shape, v, faces, material, shapes); faces.clear(); continue; }
  }
  close_shape(shape, v, faces, material, shapes);
  return true;
}

// ---- packer ------------------------------------------------------------------------------------------------------------
struct Model { std::vector<ObjShape> shapes; std::vector<ObjMaterial> materials; unsigned long size; };

clrt_material fill_material(const ObjMaterial& m) {       // Scene.cpp:97-126
  clrt_material o; memset(&o, 0, sizeof(o));
  o.ambient.x = m.ambient[0]; o.ambient.y = m.ambient[1]; o.ambient.z = m.ambient[2];
  o.diffuse.x = m.diffuse[0]; o.diffuse.y = m.diffuse[1]; o.diffuse.z = m.diffuse[2];
  o.specular.x = m.specular[0]; o.specular.y = m.specular[1]; o.specular.z = m.specular[2];
  o.emission.x = m.emission[0]; o.emission.y = m.emission[1]; o.emission.z = m.emission[2];
  o.transmittance.x = m.transmittance[0]; o.transmittance.y = m.transmittance[1]; o.transmittance.z = m.transmittance[2];
  o.dissolve = m.dissolve; o.illum = (float)m.illum; o.ior = m.ior; o.shininess = m.shininess;
  return o;
}
inline bool feq(float a, float b) { const float d = a - b; return d <= FLT_EPSILON && d >= -FLT_EPSILON; }       // RTKernelUtils.h:125-129
inline bool f3eq(const clrt_float4& a, const clrt_float4& b) { return feq(a.x, b.x) && feq(a.y, b.y) && feq(a.z, b.z); }
bool material_equals(const clrt_material& a, const clrt_material& b) {                                            // Material.h:69-81
  return f3eq(a.ambient, b.ambient) && f3eq(a.diffuse, b.diffuse) && f3eq(a.specular, b.specular) && f3eq(a.transmittance, b.transmittance) &&
         f3eq(a.emission, b.emission) && feq(a.shininess, b.shininess) && feq(a.ior, b.ior) && feq(a.dissolve, b.dissolve) && feq(a.illum, b.illum);
}

bool parse_number(const std::string& s, float& out) {      // Scene.cpp:170-183
  const float f = (float)atof(s.c_str());
  if (f == 0.0f) for (size_t i = 0; i < s.size(); i++) if (s[i] != '0' && s[i] != '.') return false;
  out = f;
  return true;
}
int string_to_floats(const std::string& str, float* result, int size) {       // Scene.cpp:185-202 (split on ' ', runs compressed)
  std::vector<std::string> parts;
  std::string cur;
  bool in_sep = false;
  parts.push_back(std::string());
  for (size_t i = 0; i < str.size(); i++) {
    if (str[i] == ' ') { if (!in_sep) { parts.push_back(std::string()); in_sep = true; } }
    else { parts.back().push_back(str[i]); in_sep = false; }
  }
  const int count = (int)parts.size() < size ? (int)parts.size() : size;
  int index = 0;
  for (int i = 0; i < count; i++) { float v = 0.0f; if (parse_number(parts[i], v)) result[index++] = v; }
  return index;
}

bool file_exists(const std::string& p) { std::ifstream f(p.c_str()); return (bool)f; }

// The reference runs on Windows: its scene files and .bat command lines use back-slashes and spell names in whatever case
// (Bin/ClumsyDragon.bat asks for ..\clumsydragon.txt, the file is ClumsyDragon.txt).  A path that does not exist as written is
// resolved component by component, ignoring case; if that fails too the path is returned unchanged and the open reports it.
std::string ci_resolve(const std::string& p) {
  if (p.empty() || file_exists(p)) return p;
  std::string cur = p[0] == '/' ? "/" : "";
  size_t i = p[0] == '/' ? 1 : 0;
  while (i <= p.size()) {
    const size_t j = p.find('/', i);
    const std::string part = p.substr(i, j == std::string::npos ? std::string::npos : j - i);
    i = j == std::string::npos ? p.size() + 1 : j + 1;
    if (part.empty()) continue;
    if (part == "." || part == "..") { cur += part + "/"; continue; }
    std::string found;
    if (DIR* d = opendir(cur.empty() ? "." : cur.c_str())) {
      std::string exact, loose;
      while (struct dirent* e = readdir(d)) {
        if (part == e->d_name) exact = e->d_name;
        else if (loose.empty() && 0 == strcasecmp(part.c_str(), e->d_name)) loose = e->d_name;
      }
      closedir(d);
      found = !exact.empty() ? exact : loose;
    }
    if (found.empty()) return p;
    cur += found + (i <= p.size() ? "/" : "");
  }
  return file_exists(cur) ? cur : p;
}
std::string host_path(std::string p) {
  for (size_t i = 0; i < p.size(); i++) if (p[i] == '\\') p[i] = '/';
  return ci_resolve(p);
}
std::string dir_of(const std::string& p) { const size_t s = p.find_last_of('/'); return s == std::string::npos ? std::string("") : p.substr(0, s + 1); }
std::string resolve_model_path(std::string p, const std::string& scene_file) {
  for (size_t i = 0; i < p.size(); i++) if (p[i] == '\\') p[i] = '/';
  while (!p.empty() && (p[p.size() - 1] == '\r' || p[p.size() - 1] == ' ')) p.erase(p.size() - 1);
  if (p.empty() || file_exists(p)) return p;
  if (p[0] == '/') return ci_resolve(p);
  const std::string base = dir_of(scene_file);
  const std::string tries[3] = {base + p, base + "Release/" + p, p};
  for (int i = 0; i < 3; i++) if (file_exists(tries[i])) return tries[i];
  for (int i = 0; i < 3; i++) { const std::string r = ci_resolve(tries[i]); if (file_exists(r)) return r; }
  return p;
}

inline float fmin2(float a, float b) { return a < b ? a : b; }     // the reference's min/max macros (CLPortability.h:147-148)
inline float fmax2(float a, float b) { return a > b ? a : b; }

Result pack_scene_file(const char* filename_as_given, char*& out, uint64_t& out_size, Errata& err) {
  std::map<std::string, std::vector<std::string> > raw;
  const std::string resolved = host_path(filename_as_given ? filename_as_given : "");
  const char* filename = resolved.c_str();
  {
    std::ifstream f(filename);
    if (!f.is_open()) { CLRT_FILL_ERRATA(err, "File: " << filename << " couldn't be opened!"); return CLRayTracer::Common::Error; }
    std::string line;
    while (std::getline(f, line)) {
      if (!line.empty() && line[line.size() - 1] == '\r') line.erase(line.size() - 1);
      const size_t sp = line.find_first_of(' ');
      if (sp == std::string::npos) continue;
      raw[line.substr(0, sp)].push_back(line.substr(sp + 1));
    }
  }
  uint64_t total = sizeof(clrt_scene_header);
  uint64_t nPrims = 0, nLights = 0, nSpheres = 0, nMeshes = 0, modelBytes = 0;
  std::vector<Model> models;
  std::vector<clrt_material> uniq;
  for (std::map<std::string, std::vector<std::string> >::iterator it = raw.begin(); it != raw.end(); ++it) {
    if (it->first == "LIGHT") { nLights = it->second.size(); total += nLights * sizeof(clrt_light); }
    else if (it->first == "SPHERE") { nSpheres = it->second.size(); total += nSpheres * 16; }
    else if (it->first == "MESH") {
      nMeshes = it->second.size();
      for (size_t i = 0; i < it->second.size(); i++) {
        Model m;
        const std::string file = resolve_model_path(it->second[i], filename);
        std::string lerr;
        if (!load_obj(file, dir_of(file), m.shapes, m.materials, lerr)) { CLRT_FILL_ERRATA(err, lerr); return CLRayTracer::Common::Error; }
        unsigned long b = 0;
        for (size_t k = 0; k < m.shapes.size(); k++) {
          if (m.shapes[k].positions.size() / 3 > 65536) {
            CLRT_FILL_ERRATA(err, "sub-mesh " << k << " of " << file << " has more than 65536 vertices (16-bit indices, SceneBufferParser.h:160)");
            return CLRayTracer::Common::Error;
          }
          b += m.shapes[k].indices.size() * 2 + m.shapes[k].positions.size() / 3 * 16 + sizeof(clrt_mesh_header);
        }
        m.size = b + sizeof(clrt_model_header);
        modelBytes += m.size;
        std::map<unsigned, unsigned> xf;                 // Scene.cpp:129-167
        for (unsigned j = 0; j < m.materials.size(); j++) {
          const clrt_material mat = fill_material(m.materials[j]);
          bool found = false;
          for (unsigned k = 0; k < uniq.size(); k++) if (material_equals(mat, uniq[k])) { xf[j] = k; found = true; break; }
          if (!found) { xf[j] = (unsigned)uniq.size(); uniq.push_back(mat); }
        }
        for (size_t k = 0; k < m.shapes.size(); k++)
          for (size_t j = 0; j < m.shapes[k].material_ids.size(); j++) m.shapes[k].material_ids[j] = (int)xf[(unsigned)m.shapes[k].material_ids[j]];
        models.push_back(m);
      }
      total += modelBytes;
    } else continue;
    nPrims += it->second.size();
    total += uniq.size() * sizeof(clrt_material);        // sic: once per recognised key (Scene.cpp:313)
  }
  char* buf = new char[total];
  memset(buf, 0, total);
  clrt_scene_header* sh = reinterpret_cast<clrt_scene_header*>(buf);
  sh->totalDataSize = total; sh->numberOfPrimitives = nPrims; sh->numberOfLights = nLights; sh->numberOfSpheres = nSpheres;
  sh->numberOfModels = models.size(); sh->modelBufferSize = modelBytes; sh->numberOfMaterials = uniq.size();
  char* lights = buf + sizeof(clrt_scene_header);
  char* spheres = lights + nLights * sizeof(clrt_light);
  char* mats = spheres + nSpheres * 16;
  char* model_base = mats + uniq.size() * sizeof(clrt_material);
  float num[4];
  if (raw.count("LIGHT")) {
    const std::vector<std::string>& v = raw["LIGHT"];
    for (size_t i = 0; i < v.size(); i++) {
      if (string_to_floats(v[i], num, 4) != 4) { CLRT_FILL_ERRATA(err, "Light array contains some invalid values"); continue; }
      memcpy(lights + i * 16, num, 16);
    }
  }
  if (raw.count("SPHERE")) {
    const std::vector<std::string>& v = raw["SPHERE"];
    for (size_t i = 0; i < v.size(); i++) {
      if (string_to_floats(v[i], num, 4) != 4) { CLRT_FILL_ERRATA(err, "Sphere array contains some invalid values"); continue; }
      memcpy(spheres + i * 16, num, 16);
    }
  }
  for (size_t i = 0; i < uniq.size(); i++) memcpy(mats + i * sizeof(clrt_material), &uniq[i], sizeof(clrt_material));
  sh->modelsBoundingBox.bounds[0].x = sh->modelsBoundingBox.bounds[0].y = sh->modelsBoundingBox.bounds[0].z = FLT_MAX;
  sh->modelsBoundingBox.bounds[1].x = sh->modelsBoundingBox.bounds[1].y = sh->modelsBoundingBox.bounds[1].z = FLT_MIN;
  uint64_t sceneTris = 0;
  char* mp = model_base;
  for (size_t i = 0; i < nMeshes && i < models.size(); i++) {
    Model& model = models[i];
    clrt_model_header* mh = reinterpret_cast<clrt_model_header*>(mp);
    mh->dataSize = model.size; mh->numberOfSubmeshes = model.shapes.size();
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {FLT_MIN, FLT_MIN, FLT_MIN};
    uint64_t modelTris = 0;
    char* sp = mp + sizeof(clrt_model_header);
    for (size_t m = 0; m < model.shapes.size(); m++) {
      const ObjShape& s = model.shapes[m];
      const unsigned nv = (unsigned)(s.positions.size() / 3), ni = (unsigned)s.indices.size();
      clrt_mesh_header hdr; memset(&hdr, 0, sizeof(hdr));
      hdr.dataSize = (uint64_t)ni * 2 + (uint64_t)nv * 16 + sizeof(clrt_mesh_header);
      hdr.numberOfVertices = nv; hdr.numberOfIndices = ni; hdr.numberOfTriangles = ni / 3;
      // Scene.cpp:411 reads material_ids[0] unconditionally; a shape whose faces all have fewer than three vertices has none (the
      // reference reads past an empty vector there): such a sub-mesh gets tinyobj's "no material" id
      hdr.materialIndex = (uint64_t)(int64_t)(s.material_ids.empty() ? -1 : s.material_ids[0]);
      memcpy(sp, &hdr, sizeof(hdr));                      // sub-mesh blocks are only 2-byte aligned: no typed stores
      modelTris += ni / 3;
      char* vp = sp + sizeof(clrt_mesh_header);
      for (unsigned vi = 0; vi < nv; vi++) {
        const float p[4] = {s.positions[3 * vi], s.positions[3 * vi + 1], s.positions[3 * vi + 2], 0.0f};
        for (int a = 0; a < 3; a++) { mn[a] = fmin2(mn[a], p[a]); mx[a] = fmax2(mx[a], p[a]); }
        memcpy(vp + (size_t)vi * 16, p, 16);
      }
      char* ip = vp + (size_t)nv * 16;
      for (unsigned ii = 0; ii < ni; ii++) { const uint16_t ix = (uint16_t)s.indices[ii]; memcpy(ip + (size_t)ii * 2, &ix, 2); }
      mh->boundingBox.bounds[0].x = mn[0]; mh->boundingBox.bounds[0].y = mn[1]; mh->boundingBox.bounds[0].z = mn[2];
      mh->boundingBox.bounds[1].x = mx[0]; mh->boundingBox.bounds[1].y = mx[1]; mh->boundingBox.bounds[1].z = mx[2];
      // merge(model box, scene box) (AABB.h:92-98): min/max(a, b) with a = model box
      clrt_aabb& sb = sh->modelsBoundingBox;
      sb.bounds[0].x = fmin2(mn[0], sb.bounds[0].x); sb.bounds[0].y = fmin2(mn[1], sb.bounds[0].y); sb.bounds[0].z = fmin2(mn[2], sb.bounds[0].z);
      sb.bounds[1].x = fmax2(mx[0], sb.bounds[1].x); sb.bounds[1].y = fmax2(mx[1], sb.bounds[1].y); sb.bounds[1].z = fmax2(mx[2], sb.bounds[1].z);
      sp += hdr.dataSize;
    }
    mh->numberOfTriangles = modelTris;
    sceneTris += modelTris;
    mp += model.size;
  }
  sh->totalNumberOfTriangles = sceneTris;
  out = buf; out_size = total;
  return CLRayTracer::Common::Success;
}

}  // namespace

namespace CLRayTracer {
Common::Result Scene::load(const char* filename, Common::Errata& err) {
  char* buf = NULL; uint64_t size = 0;
  // Result + Errata is the interface's error convention: a malformed OBJ (face index outside the vertex list -> std::out_of_range) or
  // an allocation failure is reported, not thrown through the caller
  try {
    if (pack_scene_file(filename, buf, size, err) != Common::Success) return Common::Error;
  } catch (const std::exception& e) {
    CLRT_FILL_ERRATA(err, (std::string("Scene::load: ") + e.what()).c_str()); return Common::Error;
  }
  delete[] _hostSceneData;
  _hostSceneData = buf; _sceneDataSize = size;
  if (_device) { clrt_scene_destroy(_device); _device = NULL; }
  return Common::Success;
}
}  // namespace CLRayTracer

// ---- C entry points (include/clrt_host.h) ---------------------------------------------------------------------------
static thread_local std::string g_host_error;
extern "C" const char* clrt_host_last_error(void) { return g_host_error.c_str(); }
extern "C" int clrt_host_scene_load(const char* scene_file, void** out_buf, uint64_t* out_bytes) {
  if (!scene_file || !out_buf || !out_bytes) { g_host_error = "clrt_host_scene_load: NULL argument"; return 1; }
  Errata err; char* buf = NULL; uint64_t size = 0;
  // no C++ exception may cross the C boundary (a face index outside the vertex list throws std::out_of_range, a huge scene bad_alloc)
  try {
    if (pack_scene_file(scene_file, buf, size, err) != CLRayTracer::Common::Success) { g_host_error = err.message(); return 1; }
  } catch (const std::exception& e) {
    g_host_error = std::string("clrt_host_scene_load: ") + e.what(); return 1;
  } catch (...) {
    g_host_error = "clrt_host_scene_load: unknown exception"; return 1;
  }
  *out_buf = buf; *out_bytes = size;
  return 0;
}
extern "C" void clrt_host_free(void* buf) { delete[] static_cast<char*>(buf); }

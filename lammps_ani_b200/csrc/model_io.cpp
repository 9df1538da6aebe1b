// model_io.cpp -- `.aniw` reader of libani_b200 (layout: lammps_ani_b200/aniw.py).
// Replaces torch::jit::load + module_to_dtype of the reference runner (src/ani_csrc/ani.cpp:13-33,46-61).
#include <cstring>
#include <fstream>

#include "common.cuh"

namespace anib {

namespace {
struct Reader {
  const std::vector<unsigned char>& buf;
  size_t off = 0;
  bool ok = true;
  explicit Reader(const std::vector<unsigned char>& b) : buf(b) {}
  template <typename T>
  T get() {
    T v{};
    if (off + sizeof(T) > buf.size()) { ok = false; return v; }
    std::memcpy(&v, buf.data() + off, sizeof(T));
    off += sizeof(T);
    return v;
  }
  template <typename T>
  void get_n(std::vector<T>& out, size_t n) {
    out.resize(n);
    if (off + n * sizeof(T) > buf.size()) { ok = false; return; }
    std::memcpy(out.data(), buf.data() + off, n * sizeof(T));
    off += n * sizeof(T);
  }
};
}  // namespace

bool load_aniw(const std::string& path_in, HostModel& m, std::string& err) {
  std::string path = path_in;
  auto ends_with = [](const std::string& s, const std::string& suf) {
    return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
  };
  if (ends_with(path, ".pt")) {
    // the reference's TorchScript archive: parameters must have been extracted once on the host
    std::string side = path + ".aniw";
    std::ifstream probe(side, std::ios::binary);
    if (!probe.good()) {
      err = "model file '" + path + "' is a TorchScript archive; extract it once with\n  python -m lammps_ani_b200.extract " +
            path + " " + side + "\n(libani_b200 does not link libtorch)";
      return false;
    }
    path = side;
  }
  std::ifstream f(path, std::ios::binary);
  if (!f.good()) { err = "cannot open model file '" + path + "'"; return false; }
  std::vector<unsigned char> buf((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  if (buf.size() < 80 || std::memcmp(buf.data(), "ANIW0001", 8) != 0) {
    err = "'" + path + "' is not an ANIW0001 parameter file";
    return false;
  }
  Reader r(buf);
  r.off = 8;
  m.S = r.get<int32_t>(); m.nR = r.get<int32_t>(); m.nA = r.get<int32_t>(); m.nZ = r.get<int32_t>();
  m.M = r.get<int32_t>();
  int maxl = r.get<int32_t>();
  m.use_repulsion = r.get<int32_t>();
  (void)r.get<int32_t>();
  m.Rcr = r.get<double>(); m.Rca = r.get<double>(); m.EtaR = r.get<double>(); m.EtaA = r.get<double>();
  m.Zeta = r.get<double>();
  if (m.S < 1 || m.S > kMaxSpecies || m.nR < 1 || m.nR > kMaxShfR || m.nA < 1 || m.nA > kMaxShfA || m.nZ < 1 ||
      m.nZ > kMaxShfZ || m.M < 1 || maxl < 1 || maxl > kMaxLayers) {
    err = "'" + path + "': unsupported model geometry";
    return false;
  }
  r.get_n(m.ShfR, m.nR); r.get_n(m.ShfA, m.nA); r.get_n(m.ShfZ, m.nZ);
  m.symbols.clear();
  for (int s = 0; s < m.S; s++) {
    char sym[5] = {0};
    for (int k = 0; k < 4; k++) sym[k] = (char)r.get<unsigned char>();
    m.symbols.emplace_back(sym);
  }
  r.get_n(m.sae, m.S);
  m.dims.assign(m.S, {});
  for (int s = 0; s < m.S; s++) {
    int nl = r.get<int32_t>();
    if (!r.ok || nl < 1 || nl > kMaxLayers) { err = "'" + path + "': bad layer count"; return false; }
    std::vector<int32_t> d;
    r.get_n(d, nl + 1);
    m.dims[s].assign(d.begin(), d.end());
    if (m.dims[s][0] != m.aev_len() || m.dims[s].back() != 1) {
      err = "'" + path + "': network input/output size does not match the AEV geometry";
      return false;
    }
  }
  m.W.assign(m.M, {}); m.b.assign(m.M, {});
  for (int mm = 0; mm < m.M; mm++) {
    m.W[mm].assign(m.S, {}); m.b[mm].assign(m.S, {});
    for (int s = 0; s < m.S; s++) {
      int nl = (int)m.dims[s].size() - 1;
      m.W[mm][s].assign(nl, {}); m.b[mm][s].assign(nl, {});
      for (int l = 0; l < nl; l++) {
        r.get_n(m.W[mm][s][l], (size_t)m.dims[s][l] * m.dims[s][l + 1]);
        r.get_n(m.b[mm][s][l], (size_t)m.dims[s][l + 1]);
      }
    }
  }
  if (m.use_repulsion) {
    m.rep_cutoff = r.get<double>();
    m.rep_cutoff_fn = r.get<int32_t>();
    (void)r.get<int32_t>();
    r.get_n(m.rep_y, (size_t)m.S * m.S); r.get_n(m.rep_sa, (size_t)m.S * m.S); r.get_n(m.rep_k, (size_t)m.S * m.S);
    if (!r.ok || !(m.rep_cutoff > 0.0) || m.rep_cutoff_fn < 0 || m.rep_cutoff_fn > 1) { err = "'" + path + "': bad repulsion block"; return false; }
  }
  if (!r.ok || r.off != buf.size()) { err = "'" + path + "': truncated or oversized parameter file"; return false; }
  return true;
}

}  // namespace anib

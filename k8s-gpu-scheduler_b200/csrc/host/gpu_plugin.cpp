// gpu_plugin.cpp — host-side mirror of pkg/plugins/gpu_plugin/gpu_plugins.go around the GPU
// arithmetic: snapshot builder (strings -> structure-of-arrays tables) and the plugin surface
// (Name / PreScore / Score / NormalizeScore).  Pure host C++; see include/gpusched_host.h.
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/gpusched_host.h"
#include "dcgm.h"

namespace {

std::string normalise(const std::string& s) {  // request.index.replace("-", "_")
  std::string r = s;
  for (char& c : r)
    if (c == '-') c = '_';
  return r;
}

// json.Unmarshal([]byte(uuids), &tmpUuids) for a JSON array of strings (gpu_plugins.go:541-542).
// Anything that is not such an array leaves the list empty, as the ignored Unmarshal error does.
bool parse_json_string_array(const char* s, std::vector<std::string>* out) {
  out->clear();
  if (!s) return false;
  const char* p = s;
  auto ws = [&] { while (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r') ++p; };
  ws();
  if (*p != '[') return false;
  ++p;
  ws();
  if (*p == ']') return true;
  auto at_end = [&]() {  // encoding/json validates the whole input first: trailing data -> error, nothing decoded
    ws();
    if (*p == '\0') return true;
    out->clear();
    return false;
  };
  for (;;) {
    ws();
    if (*p != '"') {
      // a scalar of another type: encoding/json records an UnmarshalTypeError, leaves the element's zero
      // value ("") and goes on; nested arrays/objects are not reproduced (list left empty)
      const char* q = p;
      while (*q && *q != ',' && *q != ']' && *q != '[' && *q != '{' && *q != '"') ++q;
      std::string tok(p, q);
      while (!tok.empty() && isspace((unsigned char)tok.back())) tok.pop_back();
      bool scalar = tok == "null" || tok == "true" || tok == "false";
      if (!scalar && !tok.empty()) {
        char* endp = nullptr;
        strtod(tok.c_str(), &endp);
        scalar = endp && *endp == '\0';
      }
      if (!scalar || (*q != ',' && *q != ']')) { out->clear(); return false; }
      out->push_back("");
      p = q;
      if (*p == ',') { ++p; continue; }
      ++p;
      return at_end();
    }
    ++p;
    std::string cur;
    // encoding/json's scanner rejects the whole document on a malformed string: an escape other
    // than \" \\ \/ \b \f \n \r \t \uXXXX, a \u without four hex digits, a raw control character
    auto hex4 = [](const char* q, unsigned* v) {  // q -> the four digits after "\u"
      *v = 0;
      for (int i = 0; i < 4; ++i) {
        const unsigned char ch = (unsigned char)q[i];  // a NUL ends the scan here: never read past it
        if (!isxdigit(ch)) return false;
        *v = *v * 16 + (unsigned)(isdigit(ch) ? ch - '0' : tolower(ch) - 'a' + 10);
      }
      return true;
    };
    auto utf8 = [&cur](unsigned cp) {
      if (cp < 0x80) {
        cur += (char)cp;
      } else if (cp < 0x800) {
        cur += (char)(0xC0 | (cp >> 6));
        cur += (char)(0x80 | (cp & 0x3F));
      } else if (cp < 0x10000) {
        cur += (char)(0xE0 | (cp >> 12));
        cur += (char)(0x80 | ((cp >> 6) & 0x3F));
        cur += (char)(0x80 | (cp & 0x3F));
      } else {
        cur += (char)(0xF0 | (cp >> 18));
        cur += (char)(0x80 | ((cp >> 12) & 0x3F));
        cur += (char)(0x80 | ((cp >> 6) & 0x3F));
        cur += (char)(0x80 | (cp & 0x3F));
      }
    };
    bool bad = false;
    while (*p && *p != '"' && !bad) {
      if ((unsigned char)*p < 0x20) {
        bad = true;
      } else if (*p == '\\') {
        ++p;
        switch (*p) {
          case '"': cur += '"'; ++p; break;
          case '\\': cur += '\\'; ++p; break;
          case '/': cur += '/'; ++p; break;
          case 'n': cur += '\n'; ++p; break;
          case 't': cur += '\t'; ++p; break;
          case 'r': cur += '\r'; ++p; break;
          case 'b': cur += '\b'; ++p; break;
          case 'f': cur += '\f'; ++p; break;
          case 'u': {  // \uXXXX -> UTF-8, surrogate pairs combined, lone surrogates U+FFFD (as Go's unquote)
            unsigned v = 0, lo = 0;
            if (!hex4(p + 1, &v)) { bad = true; break; }
            p += 5;
            if (v >= 0xD800 && v < 0xDC00 && p[0] == '\\' && p[1] == 'u' && hex4(p + 2, &lo) && lo >= 0xDC00 && lo < 0xE000) {
              v = 0x10000 + ((v - 0xD800) << 10) + (lo - 0xDC00);
              p += 6;
            } else if (v >= 0xD800 && v < 0xE000) {
              v = 0xFFFD;
            }
            utf8(v);
            break;
          }
          default: bad = true;  // includes the NUL of a document that ends inside an escape
        }
      } else {
        cur += *p++;
      }
    }
    if (bad) { out->clear(); return false; }
    if (*p != '"') { out->clear(); return false; }
    ++p;
    out->push_back(cur);
    ws();
    if (*p == ',') { ++p; continue; }
    if (*p == ']') { ++p; return at_end(); }
    out->clear();
    return false;
  }
}

struct Node {
  std::string name;
  int model = GS_MODEL_NONE;
  std::vector<std::string> uuids;
  bool slo_error = false;  // GetSLOs returned an error for this node
  std::vector<gsp_dcgm::Sample> dcgm;  // Prometheus samples of the node's DCGM exporter
};

struct PodIn {
  std::string name, ns, phase, cvd, slo;
  bool has_cvd = false, has_slo = false;
};

}  // namespace

struct gsp_builder {
  std::vector<std::string> apps;         // configurations index
  std::vector<std::string> columns;      // configurations columns
  std::vector<float> conf;               // [A x K]
  std::vector<float> interf;             // [2 x A x A]
  std::vector<int32_t> col_of;           // [2 x 32]
  std::vector<Node> nodes;
  std::unordered_map<std::string, int> node_index;
  std::vector<PodIn> pods;
  int redis_reachable = 1;
  // built tables
  bool built = false;
  int max_gpus = 1;
  std::vector<uint8_t> node_model, node_ngpu;
  std::vector<uint32_t> gpu_off, res_app;
  std::vector<float> res_slo;
  gs_cluster cluster{};
  gs_predictions preds{};
  std::string err;

  uint32_t app_of(const std::string& pod_name) const {
    const std::string req = normalise(pod_name);
    for (size_t a = 0; a < apps.size(); ++a)
      if (!apps[a].empty() && req.find(apps[a]) != std::string::npos) return (uint32_t)a;
    return GS_APP_UNKNOWN;
  }
};

struct gsp_plugin {
  gs_ctx* ctx = nullptr;
  gsp_builder* b = nullptr;
  std::vector<uint8_t> row_score, row_sel;
  bool have_row = false;
  uint32_t app = GS_APP_UNKNOWN;  // the cycle's pod, as PreScore saw it
  float slo = 0.0f;
  std::string status;
};

extern "C" {

float gsp_parse_slo(const char* value, int* ok) {
  if (ok) *ok = 0;
  if (!value || !*value) return 0.0f;  // utils.GetEnv returned "" (gpu_plugins.go:461-462)
  // strconv.ParseFloat rejects surrounding white space; strtof would skip it
  if (isspace((unsigned char)value[0])) return 0.0f;
  errno = 0;
  char* end = nullptr;
  float v = strtof(value, &end);
  if (end == value || *end != '\0') return 0.0f;                     // syntax error -> err != nil -> 0 (:466-468)
  if (errno == ERANGE && (v == HUGE_VALF || v == -HUGE_VALF)) return 0.0f;  // ErrRange -> err != nil -> 0
  if (ok) *ok = 1;
  return v;
}

gsp_builder* gsp_builder_create(void) {
  gsp_builder* b = new gsp_builder();
  b->col_of.assign(2 * (GS_MAX_GPUS_PER_NODE + 1), -1);
  return b;
}
void gsp_builder_destroy(gsp_builder* b) { delete b; }
const char* gsp_builder_last_error(const gsp_builder* b) { return b ? b->err.c_str() : "null builder"; }

int gsp_builder_set_configurations(gsp_builder* b, const char* const* index, int n_rows, const char* const* columns,
                                   int n_cols, const float* values) {
  if (!b || n_rows < 0 || n_cols < 0 || (n_rows && !index) || (n_cols && !columns) || (n_rows && n_cols && !values))
    return GS_ERR_INVALID;
  b->built = false;
  b->apps.assign(index, index + n_rows);
  b->columns.assign(columns, columns + n_cols);
  b->conf.assign(values, values + (size_t)n_rows * n_cols);
  b->col_of.assign(2 * (GS_MAX_GPUS_PER_NODE + 1), -1);
  for (int c = 0; c < n_cols; ++c) {  // "%dP_%s" (gpu_plugins.go:583, 636-643)
    const std::string& name = b->columns[c];
    size_t pos = name.find("P_");
    if (pos == std::string::npos || pos == 0) continue;
    bool digits = true;
    for (size_t i = 0; i < pos; ++i) digits = digits && isdigit((unsigned char)name[i]);
    if (!digits) continue;
    long k = strtol(name.substr(0, pos).c_str(), nullptr, 10);
    const std::string model = name.substr(pos + 2);
    int mi = model == "A30" ? 0 : model == "V100" ? 1 : -1;
    if (mi < 0 || k < 0 || k > GS_MAX_GPUS_PER_NODE) continue;
    int32_t& slot = b->col_of[mi * (GS_MAX_GPUS_PER_NODE + 1) + k];
    if (slot < 0) slot = c;
  }
  b->interf.assign((size_t)2 * n_rows * n_rows, 0.0f);
  return GS_OK;
}

int gsp_builder_set_interference(gsp_builder* b, const char* const* index, int n_rows, const char* const* columns,
                                 int n_cols, const float* values) {
  if (!b || n_rows < 0 || n_cols < 0 || (n_rows && !index) || (n_cols && !columns) || (n_rows && n_cols && !values))
    return GS_ERR_INVALID;
  b->built = false;
  const size_t A = b->apps.size();
  std::unordered_map<std::string, int> app_id;
  for (size_t a = 0; a < A; ++a) app_id.emplace(b->apps[a], (int)a);
  b->interf.assign(2 * A * A, 0.0f);
  std::vector<int> col_app(n_cols, -1);
  for (int c = 0; c < n_cols; ++c) {
    auto it = app_id.find(columns[c]);
    if (it != app_id.end()) col_app[c] = it->second;
  }
  for (int r = 0; r < n_rows; ++r) {
    const std::string key = index[r];
    int mi = -1;
    std::string app;
    if (key.size() > 4 && key.compare(key.size() - 4, 4, "_A30") == 0) {
      mi = 0;
      app = key.substr(0, key.size() - 4);
    } else if (key.size() > 5 && key.compare(key.size() - 5, 5, "_V100") == 0) {
      mi = 1;
      app = key.substr(0, key.size() - 5);
    }
    auto it = app_id.find(app);
    if (mi < 0 || it == app_id.end()) continue;
    for (int c = 0; c < n_cols; ++c)
      if (col_app[c] >= 0) b->interf[(size_t)mi * A * A + (size_t)it->second * A + col_app[c]] = values[(size_t)r * n_cols + c];
  }
  return GS_OK;
}

int gsp_builder_add_node(gsp_builder* b, const char* name, const char* redis_value) {
  if (!b || !name) return GS_ERR_INVALID;
  b->built = false;
  Node n;
  n.name = name;
  if (n.name.find("a30") != std::string::npos) n.model = GS_MODEL_A30;       // tested first (:478)
  else if (n.name.find("gpu") != std::string::npos) n.model = GS_MODEL_V100;  // (:497)
  parse_json_string_array(redis_value, &n.uuids);
  if ((int)n.uuids.size() > GS_MAX_GPUS_PER_NODE) {
    b->err = "node " + n.name + " lists more than 31 UUIDs";
    return GS_ERR_UNSUPPORTED;
  }
  b->node_index[n.name] = (int)b->nodes.size();
  b->nodes.push_back(std::move(n));
  return (int)b->nodes.size() - 1;
}

int gsp_builder_add_pod(gsp_builder* b, const char* name, const char* ns, const char* phase, const char* cvd,
                        const char* slo_env) {
  if (!b || !name) return GS_ERR_INVALID;
  b->built = false;
  PodIn p;
  p.name = name;
  p.ns = ns ? ns : "default";
  p.phase = phase ? phase : "";
  if (cvd) { p.cvd = cvd; p.has_cvd = true; }
  if (slo_env) { p.slo = slo_env; p.has_slo = true; }
  b->pods.push_back(std::move(p));
  return GS_OK;
}

void gsp_builder_set_redis_reachable(gsp_builder* b, int reachable) {
  if (b) { b->redis_reachable = reachable ? 1 : 0; b->built = false; }
}

uint32_t gsp_builder_app_of(const gsp_builder* b, const char* pod_name) {
  return (b && pod_name) ? b->app_of(pod_name) : GS_APP_UNKNOWN;
}

int gsp_builder_build(gsp_builder* b) {
  if (!b) return GS_ERR_INVALID;
  const size_t N = b->nodes.size();
  b->max_gpus = 1;
  for (const Node& n : b->nodes) b->max_gpus = std::max<int>(b->max_gpus, (int)n.uuids.size());
  const int G = b->max_gpus;
  // uuid -> (node, gpu); a UUID listed by several nodes belongs to each of them (uuidsSet is per node, :100-103)
  std::unordered_map<std::string, std::vector<std::pair<int, int>>> where;
  for (size_t n = 0; n < N; ++n)
    for (size_t g = 0; g < b->nodes[n].uuids.size(); ++g) where[b->nodes[n].uuids[g]].push_back({(int)n, (int)g});
  std::vector<std::vector<std::pair<uint32_t, float>>> slot(N * G);
  for (Node& n : b->nodes) n.slo_error = false;
  for (const PodIn& p : b->pods) {  // GetSLOs, gpu_plugins.go:104-157
    if (p.ns != "default") continue;             // podsDesc is built for pod.Namespace of the pending pod; demo uses default
    if (p.phase != "Running") continue;          // :105
    if (!p.has_cvd || p.cvd.empty()) continue;   // :131-133
    auto it = where.find(p.cvd);
    if (it == where.end()) continue;             // :135-136
    if (!p.has_slo || p.slo.empty()) continue;   // :137-140
    int ok = 0;
    float slo = gsp_parse_slo(p.slo.c_str(), &ok);
    for (auto [n, g] : it->second) {
      if (!ok) {
        // strconv.ParseFloat failed: GetSLOs returns the error, Logic returns it, Score fails (:141-144,:544-548)
        // (an out-of-range value also errors in Go; gsp_parse_slo reports both as !ok)
        b->nodes[n].slo_error = true;
        continue;
      }
      slot[(size_t)n * G + g].push_back({b->app_of(p.name), slo});
    }
  }
  b->node_model.resize(N);
  b->node_ngpu.resize(N);
  b->gpu_off.assign(N * G + 1, 0);
  b->res_app.clear();
  b->res_slo.clear();
  for (size_t n = 0; n < N; ++n) {
    b->node_model[n] = (uint8_t)b->nodes[n].model;
    b->node_ngpu[n] = (uint8_t)b->nodes[n].uuids.size();
    for (int g = 0; g < G; ++g) {
      for (auto& r : slot[n * G + g]) {
        b->res_app.push_back(r.first);
        b->res_slo.push_back(r.second);
      }
      b->gpu_off[n * G + g + 1] = (uint32_t)b->res_app.size();
    }
  }
  b->cluster.n_nodes = (int32_t)N;
  b->cluster.max_gpus = G;
  b->cluster.node_model = b->node_model.data();
  b->cluster.node_ngpu = b->node_ngpu.data();
  b->cluster.gpu_off = b->gpu_off.data();
  b->cluster.res_app = b->res_app.data();
  b->cluster.res_slo = b->res_slo.data();
  b->cluster.redis_reachable = b->redis_reachable;
  b->preds.n_apps = (int32_t)b->apps.size();
  b->preds.n_cols = (int32_t)b->columns.size();
  b->preds.conf = b->conf.data();
  if (b->interf.size() != (size_t)2 * b->apps.size() * b->apps.size()) b->interf.assign((size_t)2 * b->apps.size() * b->apps.size(), 0.0f);
  b->preds.interf = b->interf.data();
  b->preds.col_of = b->col_of.data();
  b->built = true;
  return GS_OK;
}

const gs_cluster* gsp_builder_cluster(const gsp_builder* b) { return (b && b->built) ? &b->cluster : nullptr; }
const gs_predictions* gsp_builder_predictions(const gsp_builder* b) { return (b && b->built) ? &b->preds : nullptr; }

const char* gsp_plugin_name(void) { return "GPU"; }

gsp_plugin* gsp_plugin_create(gs_ctx* ctx, gsp_builder* b) {
  if (!ctx || !b) return nullptr;
  if (!b->built && gsp_builder_build(b) != GS_OK) return nullptr;
  if (gs_upload_predictions(ctx, &b->preds) != GS_OK) return nullptr;
  if (gs_upload_cluster(ctx, &b->cluster, 0) != GS_OK) return nullptr;
  gsp_plugin* p = new gsp_plugin();
  p->ctx = ctx;
  p->b = b;
  return p;
}
void gsp_plugin_destroy(gsp_plugin* p) { delete p; }
const char* gsp_plugin_last_status(const gsp_plugin* p) { return p ? p->status.c_str() : "null plugin"; }

int gsp_plugin_prescore(gsp_plugin* p, const char* pod_name, const char* slo_env) {
  if (!p || !pod_name) return GSP_ERROR;
  const size_t N = p->b->nodes.size();
  p->row_score.assign(N, 0);
  p->row_sel.assign(N, GS_SEL_NONE);
  p->have_row = false;
  uint32_t app = p->b->app_of(pod_name);
  float slo = gsp_parse_slo(slo_env, nullptr);
  p->app = app;
  p->slo = slo;
  int rc = gs_score_batch(p->ctx, &app, &slo, nullptr, 1, GS_SCORE_RAW, p->row_score.data(), p->row_sel.data(), nullptr);
  if (rc != GS_OK) {
    p->status = std::string("Error in Logic() in Score(): ") + gs_last_error(p->ctx);  // :805
    return GSP_ERROR;
  }
  p->have_row = true;
  p->status.clear();
  return GSP_SUCCESS;
}

int gsp_plugin_score(gsp_plugin* p, const char* node_name, int64_t* out_score, char* out_uuid, int uuid_cap) {
  if (!p || !node_name || !out_score) return GSP_ERROR;
  *out_score = 0;
  if (out_uuid && uuid_cap > 0) out_uuid[0] = '\0';
  auto it = p->b->node_index.find(node_name);
  if (it == p->b->node_index.end()) {  // :798-801
    p->status = std::string("getting node \"") + node_name + "\" from Snapshot: not found";
    return GSP_ERROR;
  }
  if (!p->have_row) {
    p->status = "Error in Logic() in Score(): PreScore has not run for this cycle";
    return GSP_ERROR;
  }
  const Node& n = p->b->nodes[it->second];
  if (n.slo_error && p->b->redis_reachable) {  // GetSLOs failed -> Logic returns (0, err) (:544-548)
    p->status = "Error in Logic() in Score(): strconv.ParseFloat: parsing SLO of a resident pod: invalid syntax";
    return GSP_ERROR;
  }
  *out_score = p->row_score[it->second];
  unsigned sel = GS_SEL_GPU(p->row_sel[it->second]);
  if (out_uuid && uuid_cap > 0 && sel != GS_SEL_NONE && sel < n.uuids.size()) {
    strncpy(out_uuid, n.uuids[sel].c_str(), (size_t)uuid_cap - 1);
    out_uuid[uuid_cap - 1] = '\0';
  }
  p->status.clear();
  return GSP_SUCCESS;
}

int gsp_plugin_postbind(gsp_plugin* p, const char* node_name, char* out_uuid, int uuid_cap, char* out_mps_mem,
                        int mem_cap, char* out_mps_threads, int threads_cap) {
  auto put = [](char* dst, int cap, const char* src) {
    if (dst && cap > 0) {
      strncpy(dst, src, (size_t)cap - 1);
      dst[cap - 1] = '\0';
    }
  };
  put(out_uuid, uuid_cap, "");
  put(out_mps_mem, mem_cap, "");
  put(out_mps_threads, threads_cap, "");
  if (!p || !node_name) return -GSP_ERROR;
  auto it = p->b->node_index.find(node_name);
  if (it == p->b->node_index.end() || !p->have_row) {
    p->status = "PostBind: unknown node or no PreScore for this cycle";
    return -GSP_ERROR;
  }
  if (!p->b->redis_reachable) {  // klog.Fatal("Redis not found"), gpu_plugins.go:852
    p->status = "Redis not found";
    return -GSP_ERROR;
  }
  const Node& n = p->b->nodes[it->second];
  // cfgMap.Data[nodeName]: Logic's selectedUUID for this node (:759-768).  When Logic failed for the
  // node the key is missing and the reference keeps a RANDOM UUID of the node (:889); the first one
  // stands in for it here.
  std::string uuid;
  if (n.slo_error) {
    if (!n.uuids.empty()) uuid = n.uuids[0];
  } else {
    const unsigned sel = GS_SEL_GPU(p->row_sel[it->second]);
    if (sel != GS_SEL_NONE && sel < n.uuids.size()) uuid = n.uuids[sel];
  }
  // cfgMap.Data["MPS_"+nodeName]: the confIndex Logic recorded on V100 nodes inside its GPU loop
  // (:640-665): "1P_V100" unless a smaller-than-current prediction above the SLO is found
  if (!n.slo_error && n.model == GS_MODEL_V100 && !n.uuids.empty()) {
    int which = 1;
    float best = -1.0f;
    if (p->app != GS_APP_UNKNOWN) {
      for (int i = 0; i < 3; ++i) {
        const int col = p->b->col_of[(GS_MODEL_V100 - 1) * (GS_MAX_GPUS_PER_NODE + 1) + (1 << i)];
        if (col < 0) continue;
        const float v = p->b->conf[(size_t)p->app * p->b->columns.size() + (size_t)col];
        if (v > p->slo && (best == -1.0f || v < best)) {
          best = v;
          which = 1 << i;
        }
      }
    }
    if (which == 2) {  // strings.Contains(value, "2"), :897-899
      put(out_mps_mem, mem_cap, "0=16350MB");
      put(out_mps_threads, threads_cap, "50");
    } else if (which == 4) {  // :900-902
      put(out_mps_mem, mem_cap, "0=8175MB");
      put(out_mps_threads, threads_cap, "25");
    }
  }
  put(out_uuid, uuid_cap, uuid.c_str());
  p->status.clear();
  return uuid.empty() ? 0 : 1;  // the reference writes the three keys only when uuid != "" (:907)
}

int gsp_builder_add_dcgm_response(gsp_builder* b, const char* node_name, const char* prometheus_json) {
  if (!b || !node_name) return GS_ERR_INVALID;
  auto it = b->node_index.find(node_name);
  if (it == b->node_index.end()) {
    b->err = std::string("gsp_builder_add_dcgm_response: unknown node ") + node_name;
    return GS_ERR_INVALID;
  }
  std::string err;
  const int n = gsp_dcgm::parse_response(prometheus_json, &b->nodes[it->second].dcgm, &err);
  if (n < 0) {
    b->err = "gsp_builder_add_dcgm_response: " + err;
    return GS_ERR_INVALID;
  }
  return n;
}

int gsp_builder_dcgm_column(gsp_builder* b, const char* metric_name, float default_value, float* out, int n_slots) {
  if (!b || !metric_name || !out) return GS_ERR_INVALID;
  if (!b->built) {
    b->err = "gsp_builder_dcgm_column before gsp_builder_build";
    return GS_ERR_STATE;
  }
  if (n_slots != (int)b->nodes.size() * b->max_gpus) {
    b->err = "gsp_builder_dcgm_column: the column holds n_nodes * max_gpus values";
    return GS_ERR_INVALID;
  }
  int filled = 0;
  for (size_t n = 0; n < b->nodes.size(); ++n) {
    std::string err;
    const int k = gsp_dcgm::node_column(b->nodes[n].dcgm, b->nodes[n].uuids, metric_name, default_value,
                                        out + n * (size_t)b->max_gpus, b->max_gpus, &err);
    if (k < 0) {
      b->err = "gsp_builder_dcgm_column: node " + b->nodes[n].name + ": " + err;
      return GS_ERR_INVALID;
    }
    filled += k;
  }
  return filled;
}

int gsp_plugin_score_all(gsp_plugin* p, int64_t* out_scores, int n) {
  if (!p || !out_scores || n != (int)p->b->nodes.size() || !p->have_row) return GSP_ERROR;
  for (int i = 0; i < n; ++i)
    out_scores[i] = (p->b->nodes[i].slo_error && p->b->redis_reachable) ? -1 : (int64_t)p->row_score[i];
  return GSP_SUCCESS;
}

int gsp_plugin_normalize_score(gsp_plugin* p, int64_t* scores, int n) {
  if (!p) return GSP_ERROR;
  gs_normalize_scores(p->ctx, scores, n);  // the reference's NormalizeScore always returns nil (:840)
  return GSP_SUCCESS;
}

}  // extern "C"

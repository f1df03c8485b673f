// dcgm.cpp — the Prometheus / DCGM data contract as per-GPU columns (host side).
//
// Reference: pkg/prom/fetch_prom_metrics/prom_metrics.go:22-61 (ParseResponse: one Response
// {MetricName, Exporter, Value, GPU_I_ID, UUID} per element of data.result) and :63-118 (the five DCGM
// metrics queried per node), and pkg/plugins/gpu_plugin/gpu_plugins.go:162-236
// (GetDcgmMetricsForUUIDS: metrics[key][MetricName] = float32(ParseFloat(Value, 32)), key = the sample's
// UUID, or — on MIG nodes — the UUID that the k-th distinct sorted GPU_I_ID maps to).  In the reference
// the only consumer of these numbers is commented out (:551-554, :668-673, :725-729, with the defaults
// util = 0, fbFree = 1), so nothing here enters a score: the samples become structure-of-arrays columns
// [node * max_gpus + gpu] next to the cluster tables, ready for a caller that wants them.
#include <algorithm>
#include <cctype>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <utility>
#include <vector>

#include "dcgm.h"

namespace gsp_dcgm {

namespace {

// A small JSON reader: just enough structure to walk data.result[].{metric{...}, value[ts, "v"]}.
struct JVal {
  enum Type { Null, Bool, Num, Str, Arr, Obj } type = Null;
  std::string str;  // Str: the text; Num: the literal
  bool b = false;
  std::vector<JVal> arr;
  std::vector<std::pair<std::string, JVal>> obj;
  const JVal* get(const char* key) const {
    if (type != Obj) return nullptr;
    for (const auto& kv : obj)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
};

struct Parser {
  const char* p;
  const char* end;
  bool ok = true;
  void ws() {
    while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p;
  }
  bool lit(const char* s) {
    const size_t n = strlen(s);
    if ((size_t)(end - p) >= n && memcmp(p, s, n) == 0) {
      p += n;
      return true;
    }
    return false;
  }
  std::string string() {
    std::string out;
    if (p >= end || *p != '"') {
      ok = false;
      return out;
    }
    ++p;
    while (p < end && *p != '"') {
      if ((unsigned char)*p < 0x20) {
        ok = false;
        return out;
      }
      if (*p == '\\') {
        if (++p >= end) break;
        switch (*p) {
          case '"': out += '"'; break;
          case '\\': out += '\\'; break;
          case '/': out += '/'; break;
          case 'b': out += '\b'; break;
          case 'f': out += '\f'; break;
          case 'n': out += '\n'; break;
          case 'r': out += '\r'; break;
          case 't': out += '\t'; break;
          case 'u': {
            if (end - p < 5) {
              ok = false;
              return out;
            }
            unsigned v = 0;
            for (int i = 1; i <= 4; ++i) {
              const unsigned char ch = (unsigned char)p[i];
              if (!isxdigit(ch)) {
                ok = false;
                return out;
              }
              v = v * 16 + (unsigned)(isdigit(ch) ? ch - '0' : tolower(ch) - 'a' + 10);
            }
            p += 4;
            if (v < 0x80) {
              out += (char)v;
            } else if (v < 0x800) {
              out += (char)(0xC0 | (v >> 6));
              out += (char)(0x80 | (v & 0x3F));
            } else {
              out += (char)(0xE0 | (v >> 12));
              out += (char)(0x80 | ((v >> 6) & 0x3F));
              out += (char)(0x80 | (v & 0x3F));
            }
            break;
          }
          default: ok = false; return out;
        }
        ++p;
      } else {
        out += *p++;
      }
    }
    if (p >= end) {
      ok = false;
      return out;
    }
    ++p;  // closing quote
    return out;
  }
  JVal value(int depth = 0) {
    JVal v;
    ws();
    if (!ok || p >= end || depth > 64) {
      ok = false;
      return v;
    }
    if (*p == '{') {
      v.type = JVal::Obj;
      ++p;
      ws();
      if (p < end && *p == '}') {
        ++p;
        return v;
      }
      for (;;) {
        ws();
        std::string k = string();
        ws();
        if (!ok || p >= end || *p != ':') {
          ok = false;
          return v;
        }
        ++p;
        JVal item = value(depth + 1);
        if (!ok) return v;
        v.obj.emplace_back(std::move(k), std::move(item));
        ws();
        if (p < end && *p == ',') {
          ++p;
          continue;
        }
        if (p < end && *p == '}') {
          ++p;
          return v;
        }
        ok = false;
        return v;
      }
    }
    if (*p == '[') {
      v.type = JVal::Arr;
      ++p;
      ws();
      if (p < end && *p == ']') {
        ++p;
        return v;
      }
      for (;;) {
        JVal item = value(depth + 1);
        if (!ok) return v;
        v.arr.push_back(std::move(item));
        ws();
        if (p < end && *p == ',') {
          ++p;
          continue;
        }
        if (p < end && *p == ']') {
          ++p;
          return v;
        }
        ok = false;
        return v;
      }
    }
    if (*p == '"') {
      v.type = JVal::Str;
      v.str = string();
      return v;
    }
    if (lit("true")) {
      v.type = JVal::Bool;
      v.b = true;
      return v;
    }
    if (lit("false")) {
      v.type = JVal::Bool;
      return v;
    }
    if (lit("null")) return v;
    const char* q = p;
    while (q < end && (isdigit((unsigned char)*q) || *q == '-' || *q == '+' || *q == '.' || *q == 'e' || *q == 'E')) ++q;
    if (q == p) {
      ok = false;
      return v;
    }
    v.type = JVal::Num;
    v.str.assign(p, q);
    p = q;
    return v;
  }
};

}  // namespace

// ParseResponse (prom_metrics.go:22-61).  An empty body or an empty result list yields no samples; a
// body that is not JSON, or whose shape differs (the reference's type assertions would panic), fails.
int parse_response(const char* body, std::vector<Sample>* out, std::string* err) {
  if (!body || !*body) return 0;
  Parser ps{body, body + strlen(body)};
  const JVal root = ps.value();
  ps.ws();
  if (!ps.ok || ps.p != ps.end) {
    *err = "invalid JSON in the Prometheus response";
    return -1;
  }
  const JVal* data = root.get("data");
  const JVal* results = data ? data->get("result") : nullptr;
  if (!results || results->type != JVal::Arr) {
    *err = "Prometheus response without data.result[]";
    return -1;
  }
  int n = 0;
  for (const JVal& r : results->arr) {
    const JVal* value = r.get("value");
    const JVal* metric = r.get("metric");
    if (!value || value->type != JVal::Arr || value->arr.size() < 2 || value->arr[1].type != JVal::Str || !metric ||
        metric->type != JVal::Obj) {
      *err = "Prometheus sample without metric{} / value[ts, \"v\"]";
      return -1;
    }
    const JVal* name = metric->get("__name__");
    const JVal* pod = metric->get("pod");
    if (!name || name->type != JVal::Str || !pod || pod->type != JVal::Str) {
      *err = "Prometheus sample without metric.__name__ / metric.pod";
      return -1;
    }
    Sample s;
    s.metric = name->str;
    s.exporter = pod->str;
    s.value = value->arr[1].str;
    if (const JVal* g = metric->get("GPU_I_ID"))
      if (g->type == JVal::Str) s.gpu_i_id = g->str;
    if (const JVal* u = metric->get("UUID"))
      if (u->type == JVal::Str) s.uuid = u->str;
    out->push_back(std::move(s));
    ++n;
  }
  return n;
}

// GetDcgmMetricsForUUIDS' keying (gpu_plugins.go:193-228) for one node, then one metric as a column over
// the node's GPUs in Redis array order.  out[g] keeps `dflt` where the metric has no sample.
int node_column(const std::vector<Sample>& samples, const std::vector<std::string>& uuids, const char* metric,
                float dflt, float* out, int max_gpus, std::string* err) {
  for (int g = 0; g < max_gpus; ++g) out[g] = dflt;
  // MIG node: the k-th distinct GPU_I_ID (string order) stands for the k-th UUID of the node
  bool is_mig = false;
  for (const std::string& u : uuids) is_mig = is_mig || u.find("MIG") != std::string::npos;
  std::map<std::string, std::string> id_to_uuid;
  if (is_mig) {
    std::vector<std::string> ids;
    for (const Sample& s : samples) ids.push_back(s.gpu_i_id);
    std::sort(ids.begin(), ids.end());
    int count = -1;
    for (const std::string& id : ids) {
      if (id_to_uuid.count(id)) continue;
      ++count;
      if (count >= (int)uuids.size()) {  // the reference indexes past the slice here and panics
        *err = "more distinct GPU_I_ID values than UUIDs on the node";
        return -1;
      }
      id_to_uuid[id] = uuids[(size_t)count];
    }
  }
  int filled = 0;
  for (const Sample& s : samples) {
    if (s.metric != metric) continue;
    errno = 0;
    char* endp = nullptr;
    const float v = strtof(s.value.c_str(), &endp);  // float32(ParseFloat(Value, 32))
    if (endp == s.value.c_str() || *endp != '\0') {
      *err = "sample value is not a number: " + s.value;
      return -1;
    }
    const std::string& key = !s.gpu_i_id.empty() ? id_to_uuid[s.gpu_i_id] : s.uuid;
    for (size_t g = 0; g < uuids.size() && (int)g < max_gpus; ++g)
      if (uuids[g] == key) {
        out[g] = v;  // later samples overwrite earlier ones, as the map assignment does
        ++filled;
      }
  }
  return filled;
}

}  // namespace gsp_dcgm

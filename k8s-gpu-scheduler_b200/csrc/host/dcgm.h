// dcgm.h — Prometheus / DCGM samples as per-GPU columns (dcgm.cpp), used by gpu_plugin.cpp
#pragma once
#include <string>
#include <vector>

namespace gsp_dcgm {
struct Sample {  // metrics.Response, prom_metrics.go:14-20
  std::string metric, exporter, value, gpu_i_id, uuid;
};
int parse_response(const char* body, std::vector<Sample>* out, std::string* err);
int node_column(const std::vector<Sample>& samples, const std::vector<std::string>& uuids, const char* metric, float dflt,
                float* out, int max_gpus, std::string* err);
}  // namespace gsp_dcgm

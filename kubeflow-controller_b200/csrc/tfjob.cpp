// Placement contract of the reference, restated (see tfjob.hpp for the map to
// pkg/tensorflow/distributed.go), and its C-ABI face from include/accel_ps.h.
#include "tfjob.hpp"

#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <stdexcept>

#include "ps_common.h"

namespace tfjob {

const char kAlphanums[] = "bcdfghjklmnpqrstvwxz2456789";

const char *replica_type_name(ReplicaType t) {
  switch (t) {
    case ReplicaType::PS:
      return "PS";
    case ReplicaType::Worker:
      return "Worker";
    default:
      return "Local";
  }
}

std::string replica_type_lower(ReplicaType t) {
  std::string s = replica_type_name(t);
  std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return tolower(c); });
  return s;
}

const char *action_name(ActionType a) {
  static const char *names[] = {"ShouldAddPS",     "ShouldAddPSService", "ShouldAddWorker",
                                "ShouldAddWorkerService", "ShouldDelete", "Nothing"};
  return names[static_cast<int>(a)];
}

RandomSuffixSource::RandomSuffixSource(uint64_t seed) : state_(seed ? seed : 0x9e3779b97f4a7c15ull) {}

std::string RandomSuffixSource::next() {
  std::string s(kRandomLength, 'b');
  for (int i = 0; i < kRandomLength; ++i) {
    // xorshift64*; the reference's stream (Go math/rand) is not reproducible
    // outside Go, and nothing downstream depends on the particular characters
    state_ ^= state_ >> 12;
    state_ ^= state_ << 25;
    state_ ^= state_ >> 27;
    uint64_t r = state_ * 0x2545f4914f6cdd1dull;
    s[i] = kAlphanums[(r >> 33) % (sizeof(kAlphanums) - 1)];
  }
  return s;
}

std::string FixedSuffixSource::next() {
  if (i_ >= s_.size()) throw std::out_of_range("FixedSuffixSource exhausted");
  return s_[i_++];
}

std::string generate_name(const std::string &base, SuffixSource &src) {
  const size_t max_base = kMaxNameLength - kRandomLength;
  return (base.size() > max_base ? base.substr(0, max_base) : base) + src.next();
}

DistributedJob::DistributedJob(TFJob *job, int active_worker_pods, int active_ps_pods,
                               int worker_services, int ps_services, int succeeded_worker_pods,
                               SuffixSource *names)
    : job_(job),
      active_workers_(active_worker_pods),
      active_ps_(active_ps_pods),
      worker_services_(worker_services),
      ps_services_(ps_services),
      succeeded_workers_(succeeded_worker_pods),
      names_(names) {}

int DistributedJob::getTemplateIndex(ReplicaType typ) const {
  // the reference looks at Specs[0] then Specs[1] and nothing else; a job with
  // fewer than two specs makes the Go code panic on the index
  if (job_->specs.size() < 2) throw std::out_of_range("TFJob needs two replica specs");
  if (job_->specs[0].type == typ) return 0;
  if (job_->specs[1].type == typ) return 1;
  return -1;
}

static ReplicaSpec &spec_of(TFJob *job, int idx) {
  if (idx < 0) throw std::out_of_range("replica type not present in specs[0..1]");
  return job->specs[idx];
}

std::vector<Event> DistributedJob::Action() {
  std::vector<Event> events;
  const int expected_worker = spec_of(job_, getTemplateIndex(ReplicaType::Worker)).replicas -
                              succeeded_workers_;
  const int expected_ps = spec_of(job_, getTemplateIndex(ReplicaType::PS)).replicas;

  // services first, and only when none of that type exists yet
  if (worker_services_ == 0)
    events.push_back({ActionType::ShouldAddWorkerService, expected_worker - worker_services_});
  if (ps_services_ == 0)
    events.push_back({ActionType::ShouldAddPSService, expected_ps - ps_services_});

  if (active_workers_ < expected_worker) {
    compose();
    events.push_back({ActionType::ShouldAddWorker, expected_worker - active_workers_});
  } else if (active_workers_ == expected_worker) {
    events.push_back({ActionType::Nothing, 0});
  }
  if (active_ps_ < expected_ps)
    events.push_back({ActionType::ShouldAddPS, expected_ps - active_ps_});
  return events;
}

void DistributedJob::compose() {
  ReplicaSpec &w = spec_of(job_, getTemplateIndex(ReplicaType::Worker));
  ReplicaSpec &p = spec_of(job_, getTemplateIndex(ReplicaType::PS));
  job_->runtime_id = generate_name("", *names_);  // generateRuntimeID
  w.template_labels = getLabels(ReplicaType::Worker);
  p.template_labels = getLabels(ReplicaType::PS);
  composed_ = true;
}

std::map<std::string, std::string> DistributedJob::getLabels(ReplicaType typ) const {
  return {{"kubeflow.caicloud.io", "true"},
          {"job_type", replica_type_name(typ)},
          {"runtime_id", job_->runtime_id},
          {"tf_job_name", job_->name}};
}

std::string DistributedJob::getServiceName(ReplicaType typ, int index) const {
  return job_->name + "-" + replica_type_lower(typ) + "-" + std::to_string(index);
}

Service DistributedJob::GetService(ReplicaType typ, int index) {
  Service svc;
  svc.labels = getLabels(typ);
  svc.labels["index"] = std::to_string(index);
  const std::string key = getServiceName(typ, index);
  svc.name = generate_name(key + "-", *names_);
  service_names_[key] = svc.name;
  svc.port_name = kPortName;
  svc.port = kWorkerPort;
  return svc;
}

void DistributedJob::adoptServiceName(ReplicaType typ, int index, const std::string &name) {
  service_names_[getServiceName(typ, index)] = name;
}

static std::string join_hosts(const char *flag, const std::vector<std::string> &hosts) {
  std::string s = flag;
  for (const auto &h : hosts) s += h + ",";
  // the reference drops the last byte unconditionally: with zero replicas that
  // byte is the '=' of the flag itself
  s.pop_back();
  return s;
}

std::vector<std::string> DistributedJob::generateTFClusterSpec(ReplicaType typ, int index) const {
  auto hosts = [this](ReplicaType t) {
    std::vector<std::string> v;
    const int n = spec_of(job_, getTemplateIndex(t)).replicas;  // full count, not minus succeeded
    for (int i = 0; i < n; ++i) {
      auto it = service_names_.find(getServiceName(t, i));
      // a missing key reads as "" in Go: the host degenerates to ":2222"
      v.push_back((it == service_names_.end() ? std::string() : it->second) + ":" +
                  std::to_string(kWorkerPort));
    }
    return v;
  };
  return {join_hosts("--worker_hosts=", hosts(ReplicaType::Worker)),
          join_hosts("--ps_hosts=", hosts(ReplicaType::PS)),
          "--job_name=" + replica_type_lower(typ), "--task_index=" + std::to_string(index)};
}

ReplicaSpec *DistributedJob::GetSpec(ReplicaType typ, int index) {
  ReplicaSpec &spec = spec_of(job_, getTemplateIndex(typ));
  spec.template_labels["index"] = std::to_string(index);
  spec.container0_args = generateTFClusterSpec(typ, index);  // replaces the args wholesale
  return &spec;
}

bool IsLocalJob(const TFJob &job) {
  return !job.specs.empty() && job.specs[0].type == ReplicaType::Local;
}

LocalJob::LocalJob(TFJob *job, int active_pods, int succeeded, SuffixSource *names)
    : job_(job), has_pod_(active_pods == 1), succeeded_(succeeded), names_(names) {}

Event LocalJob::Action() {
  // had a succeeded result: no new pod
  if (succeeded_ > 0) return {ActionType::Nothing, 0};
  const int active = has_pod_ ? 1 : 0;
  if (active < kExpectedLocalWorkerNumber) {
    compose();
    return {ActionType::ShouldAddWorker, kExpectedLocalWorkerNumber - active};
  }
  return {ActionType::Nothing, 0};
}

void LocalJob::compose() {
  job_->runtime_id = generate_name("", *names_);  // generateRuntimeID
  spec_of(job_, 0).template_labels = labels();
  composed_ = true;
}

std::map<std::string, std::string> LocalJob::labels() const {
  return {{"kubeflow.caicloud.io", "true"},
          {"job_type", replica_type_name(job_->specs[0].type)},
          {"runtime_id", job_->runtime_id},
          {"tf_job_name", job_->name}};
}

static std::vector<std::string> split_commas(const std::string &s) {
  std::vector<std::string> out;
  size_t start = 0;
  while (true) {
    size_t c = s.find(',', start);
    out.push_back(s.substr(start, c == std::string::npos ? c : c - start));
    if (c == std::string::npos) break;
    start = c + 1;
  }
  return out;  // like Python's str.split(","): "" -> [""]
}

bool parse_cluster_argv(const std::vector<std::string> &argv, ClusterView *out, std::string *err) {
  bool have_job = false, have_idx = false;
  ClusterView v;
  v.job = ReplicaType::Worker;
  for (const auto &a : argv) {
    auto value_of = [&a](const char *flag, std::string *val) {
      const size_t n = strlen(flag);
      if (a.compare(0, n, flag) != 0) return false;
      *val = a.substr(n);
      return true;
    };
    std::string val;
    if (value_of("--worker_hosts=", &val)) {
      v.worker_hosts = split_commas(val);
    } else if (value_of("--ps_hosts=", &val)) {
      v.ps_hosts = split_commas(val);
    } else if (value_of("--job_name=", &val)) {
      if (val == "ps") {
        v.job = ReplicaType::PS;
      } else if (val == "worker") {
        v.job = ReplicaType::Worker;
      } else {
        if (err) *err = "job_name must be \"ps\" or \"worker\", got \"" + val + "\"";
        return false;
      }
      have_job = true;
    } else if (value_of("--task_index=", &val)) {
      char *end = nullptr;
      long t = strtol(val.c_str(), &end, 10);
      if (val.empty() || *end != '\0' || t < 0) {
        if (err) *err = "task_index must be a non-negative integer, got \"" + val + "\"";
        return false;
      }
      v.task_index = static_cast<int>(t);
      have_idx = true;
    }
  }
  // mnist_replica.py:99-102: both are mandatory
  if (!have_job) {
    if (err) *err = "Must specify an explicit `job_name`";
    return false;
  }
  if (!have_idx) {
    if (err) *err = "Must specify an explicit `task_index`";
    return false;
  }
  *out = v;
  return true;
}

}  // namespace tfjob

// ---- C ABI ------------------------------------------------------------------
using accel::fail;

static bool to_type(int replica_type, tfjob::ReplicaType *t) {
  if (replica_type == ACCEL_REPLICA_PS) {
    *t = tfjob::ReplicaType::PS;
  } else if (replica_type == ACCEL_REPLICA_WORKER) {
    *t = tfjob::ReplicaType::Worker;
  } else {
    return false;
  }
  return true;
}

static tfjob::TFJob two_spec_job(const char *name, int ps, int workers) {
  tfjob::TFJob j;
  j.name = name ? name : "";
  j.specs = {{ps, tfjob::ReplicaType::PS, {}, {}}, {workers, tfjob::ReplicaType::Worker, {}, {}}};
  return j;
}

static int copy_out(const std::string &s, char *buf, size_t buflen) {
  if (!buf || buflen < s.size() + 1)
    return fail(ACCEL_ERR_BUFSIZE, "output needs %zu bytes, buffer has %zu", s.size() + 1, buflen);
  memcpy(buf, s.data(), s.size());
  buf[s.size()] = '\0';
  return static_cast<int>(s.size());
}

extern "C" {

int accel_tfjob_action(int worker_replicas, int ps_replicas, int succeeded_workers,
                       int worker_services, int ps_services, int active_workers, int active_ps,
                       accel_event *out, int *n_out, int *composed) {
  if (!out || !n_out) return fail(ACCEL_ERR_INVALID, "out/n_out is NULL");
  tfjob::TFJob job = two_spec_job("job", ps_replicas, worker_replicas);
  tfjob::RandomSuffixSource names(1);
  tfjob::DistributedJob dj(&job, active_workers, active_ps, worker_services, ps_services,
                           succeeded_workers, &names);
  static const int to_abi[] = {ACCEL_ACTION_ADD_PS,     ACCEL_ACTION_ADD_PS_SERVICE,
                               ACCEL_ACTION_ADD_WORKER, ACCEL_ACTION_ADD_WORKER_SERVICE,
                               ACCEL_ACTION_DELETE,     ACCEL_ACTION_NOTHING};
  std::vector<tfjob::Event> ev = dj.Action();
  for (size_t i = 0; i < ev.size(); ++i) {
    out[i].action = to_abi[static_cast<int>(ev[i].action)];
    out[i].number = ev[i].number;
  }
  *n_out = static_cast<int>(ev.size());
  if (composed) *composed = dj.composed() ? 1 : 0;
  return ACCEL_OK;
}

int accel_tfjob_is_local_job(const int *replica_types, int n) {
  if (!replica_types || n < 1) return fail(ACCEL_ERR_INVALID, "a TFJob needs at least one spec");
  tfjob::TFJob job;
  for (int i = 0; i < n; ++i) {
    tfjob::ReplicaType t;
    if (replica_types[i] == ACCEL_REPLICA_LOCAL) {
      t = tfjob::ReplicaType::Local;
    } else if (!to_type(replica_types[i], &t)) {
      return fail(ACCEL_ERR_INVALID, "unknown replica_type %d", replica_types[i]);
    }
    job.specs.push_back({1, t, {}, {}});
  }
  return tfjob::IsLocalJob(job) ? 1 : 0;
}

int accel_tfjob_local_action(const char *job_name, const char *runtime_id_in, int first_spec_type,
                             int active_pods, int succeeded, const char *suffix,
                             accel_event *out_event, int *composed, char *buf, size_t buflen) {
  if (!job_name || !runtime_id_in || !out_event)
    return fail(ACCEL_ERR_INVALID, "job_name/runtime_id_in/out_event is NULL");
  if (active_pods < 0 || succeeded < 0)
    return fail(ACCEL_ERR_INVALID, "active_pods/succeeded must be >= 0");
  tfjob::ReplicaType t;
  if (first_spec_type == ACCEL_REPLICA_LOCAL) {
    t = tfjob::ReplicaType::Local;
  } else if (!to_type(first_spec_type, &t)) {
    return fail(ACCEL_ERR_INVALID, "unknown replica_type %d", first_spec_type);
  }
  tfjob::TFJob job;
  job.name = job_name;
  job.runtime_id = runtime_id_in;
  job.specs.push_back({1, t, {}, {}});
  try {
    tfjob::FixedSuffixSource names({suffix ? suffix : ""});
    tfjob::LocalJob lj(&job, active_pods, succeeded, &names);
    const tfjob::Event e = lj.Action();
    out_event->action = static_cast<int32_t>(e.action);
    out_event->number = e.number;
    if (composed) *composed = lj.composed() ? 1 : 0;
    std::string s = "runtime_id\t" + job.runtime_id + "\n";
    for (const auto &kv : job.specs[0].template_labels) s += kv.first + "=" + kv.second + "\n";
    return copy_out(s, buf, buflen);
  } catch (const std::exception &ex) {
    return fail(ACCEL_ERR_INVALID, "%s", ex.what());
  }
}

int accel_tfjob_sync_plan(const char *job_name, const char *runtime_id_in, int worker_replicas,
                          int ps_replicas, int succeeded_workers, int worker_services,
                          int ps_services, int active_workers, int active_ps,
                          const char *const *suffixes, int n_suffixes, char *buf, size_t buflen) {
  if (!job_name) return fail(ACCEL_ERR_INVALID, "job_name is NULL");
  if (n_suffixes < 0 || (n_suffixes > 0 && !suffixes))
    return fail(ACCEL_ERR_INVALID, "suffixes is NULL");
  std::vector<std::string> sfx;
  for (int i = 0; i < n_suffixes; ++i) {
    if (!suffixes[i] || strlen(suffixes[i]) != tfjob::kRandomLength)
      return fail(ACCEL_ERR_INVALID, "suffix %d must have exactly %d characters", i,
                  tfjob::kRandomLength);
    sfx.push_back(suffixes[i]);
  }
  auto fmt_labels = [](const std::map<std::string, std::string> &m) {
    std::string s;
    for (const auto &kv : m) s += (s.empty() ? "" : ",") + kv.first + "=" + kv.second;
    return s;
  };
  auto join = [](const std::vector<std::string> &v, const char *sep) {
    std::string s;
    for (size_t i = 0; i < v.size(); ++i) s += (i ? sep : "") + v[i];
    return s;
  };
  std::string out;
  try {
    tfjob::TFJob job = two_spec_job(job_name, ps_replicas, worker_replicas);
    job.runtime_id = runtime_id_in ? runtime_id_in : "";
    tfjob::FixedSuffixSource names(sfx);
    tfjob::DistributedJob dj(&job, active_workers, active_ps, worker_services, ps_services,
                             succeeded_workers, &names);
    const std::vector<tfjob::Event> events = dj.Action();   // compose() happens in here
    out += "runtime_id\t" + job.runtime_id + "\n";
    for (const auto &e : events)
      out += std::string("event\t") + tfjob::action_name(e.action) + "\t" +
             std::to_string(e.number) + "\n";
    for (const auto &e : events) {
      tfjob::ReplicaType t = tfjob::ReplicaType::Worker;
      bool service = false, pod = false;
      switch (e.action) {
        case tfjob::ActionType::ShouldAddWorkerService: service = true; break;
        case tfjob::ActionType::ShouldAddPSService: service = true; t = tfjob::ReplicaType::PS; break;
        case tfjob::ActionType::ShouldAddWorker: pod = true; break;
        case tfjob::ActionType::ShouldAddPS: pod = true; t = tfjob::ReplicaType::PS; break;
        default: break;
      }
      for (int i = 0; i < e.number; ++i) {
        if (service) {
          const tfjob::Service svc = dj.GetService(t, i);
          out += "service\t" + svc.name + "\t" + fmt_labels(svc.labels) + "\n";
        } else if (pod) {
          const tfjob::ReplicaSpec *spec = dj.GetSpec(t, i);   // the pod copies the template NOW
          out += std::string("pod\t") + tfjob::replica_type_name(t) + "\t" + std::to_string(i) +
                 "\t" + fmt_labels(spec->template_labels) + "\t" +
                 join(spec->container0_args, "\t") + "\n";
        }
      }
    }
    for (const auto &spec : job.specs)
      out += std::string("template\t") + tfjob::replica_type_name(spec.type) + "\t" +
             fmt_labels(spec.template_labels) + "\t" + join(spec.container0_args, " ") + "\n";
  } catch (const std::exception &ex) {
    // the Go code would panic here (index out of range) or block on its PRNG; report instead
    return fail(ACCEL_ERR_INVALID, "%s", ex.what());
  }
  return copy_out(out, buf, buflen);
}

int accel_tfjob_service_key(const char *job_name, int replica_type, int index, char *buf,
                            size_t buflen) {
  tfjob::ReplicaType t;
  if (!job_name) return fail(ACCEL_ERR_INVALID, "job_name is NULL");
  if (!to_type(replica_type, &t)) return fail(ACCEL_ERR_INVALID, "unknown replica_type %d", replica_type);
  tfjob::TFJob job = two_spec_job(job_name, 0, 0);
  tfjob::DistributedJob dj(&job, 0, 0, 0, 0, 0, nullptr);
  return copy_out(dj.getServiceName(t, index), buf, buflen);
}

int accel_tfjob_cluster_spec(int worker_replicas, int ps_replicas,
                             const char *const *worker_services, const char *const *ps_services,
                             int replica_type, int index, char *buf, size_t buflen) {
  tfjob::ReplicaType t;
  if (!to_type(replica_type, &t)) return fail(ACCEL_ERR_INVALID, "unknown replica_type %d", replica_type);
  if (worker_replicas < 0 || ps_replicas < 0)
    return fail(ACCEL_ERR_INVALID, "replica counts must be non-negative");
  // Fill serviceNames the way one sync of the controller does — GetService for
  // every worker, then every PS (pkg/controller/controller.go:376-395) — with the
  // caller's names standing in for the generated ones; an absent name leaves the
  // map without that key, as on a later sync of the reference.
  tfjob::TFJob job = two_spec_job("", ps_replicas, worker_replicas);
  tfjob::DistributedJob dj(&job, 0, 0, 0, 0, 0, nullptr);
  for (int i = 0; i < worker_replicas; ++i)
    if (worker_services && worker_services[i] && worker_services[i][0])
      dj.adoptServiceName(tfjob::ReplicaType::Worker, i, worker_services[i]);
  for (int i = 0; i < ps_replicas; ++i)
    if (ps_services && ps_services[i] && ps_services[i][0])
      dj.adoptServiceName(tfjob::ReplicaType::PS, i, ps_services[i]);
  const std::vector<std::string> argv = dj.generateTFClusterSpec(t, index);
  std::string flat;
  for (const auto &a : argv) {
    flat += a;
    flat.push_back('\0');
  }
  if (!buf || buflen < flat.size())
    return fail(ACCEL_ERR_BUFSIZE, "output needs %zu bytes, buffer has %zu", flat.size(), buflen);
  memcpy(buf, flat.data(), flat.size());
  return static_cast<int>(flat.size());
}

int accel_tfjob_labels(const char *job_name, int replica_type, const char *runtime_id, int index,
                       char *buf, size_t buflen) {
  tfjob::ReplicaType t;
  if (!job_name || !runtime_id) return fail(ACCEL_ERR_INVALID, "job_name/runtime_id is NULL");
  if (!to_type(replica_type, &t)) return fail(ACCEL_ERR_INVALID, "unknown replica_type %d", replica_type);
  tfjob::TFJob job = two_spec_job(job_name, 0, 0);
  job.runtime_id = runtime_id;
  tfjob::DistributedJob dj(&job, 0, 0, 0, 0, 0, nullptr);
  auto labels = dj.getLabels(t);
  if (index >= 0) labels["index"] = std::to_string(index);
  std::string s;
  for (const auto &kv : labels) s += kv.first + "=" + kv.second + "\n";  // std::map: sorted by key
  return copy_out(s, buf, buflen);
}

int accel_tfjob_parse_argv(const char *const *argv, int argc, int *replica_type, int *task_index,
                           int *n_workers, int *n_ps) {
  if (!argv || argc < 0) return fail(ACCEL_ERR_INVALID, "argv is NULL");
  std::vector<std::string> v;
  for (int i = 0; i < argc; ++i) v.push_back(argv[i] ? argv[i] : "");
  tfjob::ClusterView cv;
  std::string err;
  if (!tfjob::parse_cluster_argv(v, &cv, &err)) return fail(ACCEL_ERR_INVALID, "%s", err.c_str());
  if (replica_type) *replica_type = cv.job == tfjob::ReplicaType::PS ? ACCEL_REPLICA_PS : ACCEL_REPLICA_WORKER;
  if (task_index) *task_index = cv.task_index;
  if (n_workers) *n_workers = static_cast<int>(cv.worker_hosts.size());
  if (n_ps) *n_ps = static_cast<int>(cv.ps_hosts.size());
  return ACCEL_OK;
}

int accel_plan_flat(uint64_t n_elems, int n_ps, uint64_t align_elems, accel_shard *out,
                    uint64_t *padded_elems) {
  if (!out) return fail(ACCEL_ERR_INVALID, "out is NULL");
  if (n_ps < 1 || n_ps > ACCEL_MAX_PS)
    return fail(ACCEL_ERR_INVALID, "n_ps %d out of range [1,%d]", n_ps, ACCEL_MAX_PS);
  if (align_elems == 0) align_elems = 1;
  const uint64_t unit = align_elems * static_cast<uint64_t>(n_ps);
  const uint64_t padded = (n_elems + unit - 1) / unit * unit;
  const uint64_t per = padded / n_ps;
  for (int r = 0; r < n_ps; ++r) {
    out[r].offset = per * r;
    out[r].count = per;
  }
  if (padded_elems) *padded_elems = padded;
  return ACCEL_OK;
}

int accel_plan_round_robin(const uint64_t *var_elems, int n_vars, int n_ps, int first_task,
                           uint64_t align_elems, int *out_task, uint64_t *out_offset,
                           uint64_t *bucket_elems) {
  if (!var_elems || !out_task || !out_offset || !bucket_elems)
    return fail(ACCEL_ERR_INVALID, "NULL argument");
  if (n_vars < 0) return fail(ACCEL_ERR_INVALID, "n_vars is negative");
  if (n_ps < 1 || n_ps > ACCEL_MAX_PS)
    return fail(ACCEL_ERR_INVALID, "n_ps %d out of range [1,%d]", n_ps, ACCEL_MAX_PS);
  if (first_task < 0) return fail(ACCEL_ERR_INVALID, "first_task is negative");
  if (align_elems == 0) align_elems = 1;
  for (int r = 0; r < n_ps; ++r) bucket_elems[r] = 0;
  for (int k = 0; k < n_vars; ++k) {
    const int t = (first_task + k) % n_ps;  // _RoundRobinStrategy: next task per variable
    const uint64_t start = (bucket_elems[t] + align_elems - 1) / align_elems * align_elems;
    out_task[k] = t;
    out_offset[k] = start;
    bucket_elems[t] = start + var_elems[k];
  }
  return ACCEL_OK;
}

}  // extern "C"

// Host-side mirror of the reference's placement contract, in C++ because the
// reference is compiled Go and no Go toolchain exists in this image.
//
// Same names, argument meaning and observable behaviour as
// pkg/tensorflow/distributed.go (DistributedJob: Action :56-114, GetSpec
// :117-125, generateTFClusterSpec :127-159, getServiceName :161-163, GetService
// :165-188, getLabels :221-228, compose :210-219) and pkg/tensorflow/util.go
// (generateName :24-26 -> vendor/k8s.io/kubernetes/pkg/api/v1/generate.go:60-65).
// It models only what decides placement: replica counts, names, labels, argv.
// Kubernetes objects themselves (Pods, Services, owner refs) are out of scope.
#pragma once

#include <stdint.h>

#include <map>
#include <string>
#include <vector>

namespace tfjob {

enum class ReplicaType { PS, Worker, Local };

const char *replica_type_name(ReplicaType t);           // "PS" | "Worker" | "Local" (CRD spelling)
std::string replica_type_lower(ReplicaType t);          // strings.ToLower(string(typ))

enum class ActionType {  // pkg/tensorflow/types.go:28-34, same order
  ShouldAddPS,
  ShouldAddPSService,
  ShouldAddWorker,
  ShouldAddWorkerService,
  ShouldDelete,
  Nothing
};
const char *action_name(ActionType a);

struct Event {
  ActionType action;
  int number;
};

constexpr const char *kPortName = "kubeflow-port";  // distributed.go:27
constexpr int kWorkerPort = 2222;                   // distributed.go:28
constexpr int kMaxNameLength = 63;                  // generate.go:55
constexpr int kRandomLength = 5;                    // generate.go:56
extern const char kAlphanums[];                     // rand.go:77, 27 characters

// Source of the 5-character suffixes. The reference draws them from a global
// seeded PRNG (vendor/k8s.io/apimachinery/pkg/util/rand/rand.go:30-35,80-86);
// here the caller supplies the source so tests can pin the suffixes.
class SuffixSource {
 public:
  virtual ~SuffixSource() {}
  virtual std::string next() = 0;  // exactly kRandomLength characters of kAlphanums
};

class RandomSuffixSource : public SuffixSource {
 public:
  explicit RandomSuffixSource(uint64_t seed);
  std::string next() override;

 private:
  uint64_t state_;
};

class FixedSuffixSource : public SuffixSource {
 public:
  explicit FixedSuffixSource(std::vector<std::string> suffixes) : s_(std::move(suffixes)) {}
  std::string next() override;

 private:
  std::vector<std::string> s_;
  size_t i_ = 0;
};

// simpleNameGenerator.GenerateName: truncate base to 58 bytes, append suffix.
std::string generate_name(const std::string &base, SuffixSource &src);

struct ReplicaSpec {  // the fields of TFReplicaSpec that placement reads
  int replicas;
  ReplicaType type;
  std::map<std::string, std::string> template_labels;
  std::vector<std::string> container0_args;
};

struct TFJob {
  std::string name;
  std::string runtime_id;
  std::vector<ReplicaSpec> specs;  // only specs[0], specs[1] are ever inspected
};

struct Service {
  std::string name;
  std::map<std::string, std::string> labels;    // == selector
  std::string port_name;
  int port;
};

class DistributedJob {
 public:
  DistributedJob(TFJob *job, int active_worker_pods, int active_ps_pods, int worker_services,
                 int ps_services, int succeeded_worker_pods, SuffixSource *names);

  std::vector<Event> Action();
  // Mutates the shared template in place, like the reference, and returns it.
  ReplicaSpec *GetSpec(ReplicaType typ, int index);
  Service GetService(ReplicaType typ, int index);
  // Record an already-generated Service name under its key: the side effect of
  // GetService (distributed.go:171) without drawing a new suffix.
  void adoptServiceName(ReplicaType typ, int index, const std::string &name);

  std::vector<std::string> generateTFClusterSpec(ReplicaType typ, int index) const;
  std::string getServiceName(ReplicaType typ, int index) const;
  std::map<std::string, std::string> getLabels(ReplicaType typ) const;
  int getTemplateIndex(ReplicaType typ) const;  // -1 when the type is in neither spec
  bool composed() const { return composed_; }

 private:
  void compose();
  TFJob *job_;
  int active_workers_, active_ps_, worker_services_, ps_services_, succeeded_workers_;
  SuffixSource *names_;
  std::map<std::string, std::string> service_names_;
  bool composed_ = false;
};

// checker.IsLocalJob (pkg/checker/checker.go:8-14)
bool IsLocalJob(const TFJob &job);

// pkg/tensorflow/local.go: the single-replica job. NewLocalJob (:34-48) keeps the pod only when
// there is exactly ONE active pod — with two or more it logs and keeps none, so Action then asks
// for a new worker; that behaviour is kept. Action :50-74, compose :78-81, labels :83-90.
constexpr int kExpectedLocalWorkerNumber = 1;   // local.go:22-24
class LocalJob {
 public:
  LocalJob(TFJob *job, int active_pods, int succeeded, SuffixSource *names);
  Event Action();
  std::map<std::string, std::string> labels() const;
  bool composed() const { return composed_; }

 private:
  void compose();
  TFJob *job_;
  bool has_pod_;
  int succeeded_;
  SuffixSource *names_;
  bool composed_ = false;
};

// The cluster a replica sees after parsing its four flags
// (examples/workdir/mnist_replica.py:107-115).
struct ClusterView {
  ReplicaType job;
  int task_index = -1;
  std::vector<std::string> worker_hosts, ps_hosts;
};
bool parse_cluster_argv(const std::vector<std::string> &argv, ClusterView *out, std::string *err);

}  // namespace tfjob

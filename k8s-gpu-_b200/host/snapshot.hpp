// snapshot.hpp -- snapshot ingestion: the step right before the GPU
// path.  Turns what an operator holds -- AzureVmPool manifests (YAML, README.md:267-286), an ARM
// VM list (JSON, the shape `az vm list` prints) and Volcano job specs (YAML,
// GPU调度平台搭建.md:642-673) -- into the u32 structure-of-arrays tables of include/k8sgpu.h:
// tag strings, vmSize and location are dictionary-interned to ids on the host.
//
// The parsers accept the subset of YAML / JSON those documents use (block maps, block sequences,
// plain and quoted scalars; full JSON).  Nothing here runs on the GPU.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "arm_client.hpp"
#include "azurevmpool_types.hpp"

namespace k8sgpu_host {

// Generic document tree shared by the YAML-subset and JSON parsers.
struct Doc {
    enum Kind { Null, Scalar, Map, Seq } kind = Null;
    std::string scalar;
    std::vector<std::pair<std::string, Doc>> map;  // insertion ordered
    std::vector<Doc> seq;
    const Doc *get(const std::string &key) const;
    std::string str(const std::string &key, const std::string &dflt = "") const;
    long num(const std::string &key, long dflt = 0) const;
};

struct ParseError {
    std::string msg;
    int line = 0;
};

// Each returns false and fills *err on malformed input.
bool ParseYaml(const std::string &text, std::vector<Doc> *docs, ParseError *err);  // '---' separates documents
bool ParseJson(const std::string &text, Doc *doc, ParseError *err);

// String -> dense u32 id, in first-seen order.
class Interner {
  public:
    uint32_t Id(const std::string &s);
    bool Find(const std::string &s, uint32_t *id) const;
    const std::string &Str(uint32_t id) const { return strs_[id]; }
    uint32_t Size() const { return static_cast<uint32_t>(strs_.size()); }

  private:
    std::map<std::string, uint32_t> ids_;
    std::vector<std::string> strs_;
};

// What one VM size offers.  Built-ins: Standard_NC4as_T4_v3 (README.md:100,276)
// and Standard_ND96isr_H100_v5; others can be registered.
struct SkuShape {
    uint32_t sku_id, cap_gpu, cap_mcpu, cap_mib;
};

struct ReconcileTables {  // k8sgpu_pool_table + k8sgpu_vm_table, owned
    std::vector<int32_t> desired;
    std::vector<uint32_t> pool_owner, vm_owner, vm_flags, vm_ordinal;
    uint32_t n_owner_ids = 0;
    std::vector<std::string> pool_key;  // "<namespace>/<name>" per pool row
    std::vector<std::string> vm_name;   // per VM row
    // Pools left out of the table because another pool of the same resource group already claims
    // their owner tag ("<namespace>-<name>", README.md:238, is not injective): "<ns>/<name>" of the
    // dropped pool and of the pool that keeps the tag.  Never silently: callers must report these.
    std::vector<std::pair<std::string, std::string>> conflicts;
};

struct NodeTable {  // k8sgpu_node_table, owned (one node per VM)
    std::vector<uint32_t> free_gpu, free_mcpu, free_mib, cap_gpu, sku_id, zone_id, taint_mask, flags;
    std::vector<std::string> name;
};

struct PodTable {  // k8sgpu_pod_table, owned
    std::vector<uint32_t> req_gpu, req_mcpu, req_mib, sku_mask, zone_mask, tol_mask, pref_zone, priority;
    std::vector<std::string> name;
};

class SnapshotBuilder {
  public:
    SnapshotBuilder();
    void RegisterSku(const std::string &vm_size, uint32_t cap_gpu, uint32_t cap_mcpu, uint32_t cap_mib);

    // AzureVmPool manifests (one or more YAML documents).  Missing metadata.namespace = "default".
    bool AddPoolsYaml(const std::string &yaml, ParseError *err);
    // ARM VM list: JSON array of {name, resourceGroup, location, zones:[..], hardwareProfile:{vmSize}, tags:{..}}.
    bool AddVmsJson(const std::string &json, ParseError *err);
    // Volcano Job manifests: every task contributes `replicas` pods with the container limits
    // nvidia.com/gpu, cpu, memory; spec.nodeSelector "topology.kubernetes.io/zone" pins the zone,
    // "node.kubernetes.io/instance-type" pins the SKU; tolerations with key "gpu-taint-<bit>" set
    // tolerated taint bits.  A `singleInstanceType: gpu-1x-16c-32g-1gpu` template
    // (GPU调度平台搭建.md:534) is accepted too.
    bool AddJobsYaml(const std::string &yaml, ParseError *err);

    const std::vector<computev1alpha1::AzureVmPool> &Pools() const { return pools_; }
    const std::vector<std::pair<std::string, VM>> &Vms() const { return vms_; }  // (resource group, VM)

    ReconcileTables BuildReconcileTables() const;
    NodeTable BuildNodeTable();
    PodTable BuildPodTable();

    static bool ParseInstanceType(const std::string &s, uint32_t *gpus, uint32_t *mcpu, uint32_t *mib);
    static bool ParseCpu(const std::string &s, uint32_t *mcpu);     // "16", "16000m", "0.5"
    static bool ParseMemory(const std::string &s, uint32_t *mib);   // "32Gi", "512Mi", "32G"

  private:
    struct VmExtra {
        std::string vm_size, location, zone;
    };
    struct PodSpec {
        std::string name;
        uint32_t gpu = 0, mcpu = 0, mib = 0;
        std::string zone, instance_type;
        uint32_t tol = 0;
    };
    std::vector<computev1alpha1::AzureVmPool> pools_;
    std::vector<std::pair<std::string, VM>> vms_;
    std::vector<VmExtra> extra_;
    std::vector<PodSpec> pods_;
    std::map<std::string, SkuShape> skus_;
    Interner zones_;
};

}  // namespace k8sgpu_host

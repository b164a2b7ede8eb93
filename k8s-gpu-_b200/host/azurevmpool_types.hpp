// api/v1alpha1 types of the AzureVmPool CRD, mirrored field for field from the reference's Go
// structs (README.md:83-156) so that a controller written against them reads the same.
// JSON tag names are kept as comments; nothing here touches the GPU.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace computev1alpha1 {

// ImageReferenceSpec, README.md:113-118
struct ImageReferenceSpec {
    std::string Publisher;  // json:"publisher"
    std::string Offer;      // json:"offer"
    std::string SKU;        // json:"sku"
    std::string Version;    // json:"version"
};

// AzureVmPoolSpec, README.md:92-110
struct AzureVmPoolSpec {
    int32_t Replicas = 0;               // json:"replicas"  +kubebuilder:validation:Minimum=0
    std::string ResourceGroupName;      // json:"resourceGroupName"
    std::string Location;               // json:"location"
    std::string VMSize;                 // json:"vmSize"   e.g. "Standard_NC4as_T4_v3"
    std::string VnetName;               // json:"vnetName"
    std::string SubnetName;             // json:"subnetName"
    ImageReferenceSpec ImageReference;  // json:"imageReference"
    std::string AzureCredentialSecret;  // json:"azureCredentialSecret"
};

// metav1.Condition (only the fields README.md:127,310 relies on)
struct Condition {
    std::string Type;     // Provisioning / Ready / Deleting / Failed (README.md:310)
    std::string Status;   // "True" / "False"
    std::string Reason;
    std::string Message;
};

// AzureVmPoolStatus, README.md:121-128
struct AzureVmPoolStatus {
    int32_t ReadyReplicas = 0;          // json:"readyReplicas,omitempty"
    std::vector<std::string> VMs;       // json:"vms,omitempty"
    std::vector<Condition> Conditions;  // json:"conditions,omitempty"
};

struct ObjectMeta {
    std::string Namespace;
    std::string Name;
};

// AzureVmPool, README.md:136-142
struct AzureVmPool {
    ObjectMeta Metadata;
    AzureVmPoolSpec Spec;
    AzureVmPoolStatus Status;
};

// AzureVmPoolList, README.md:147-151
struct AzureVmPoolList {
    std::vector<AzureVmPool> Items;
};

}  // namespace computev1alpha1

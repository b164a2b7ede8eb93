// UNCOMPILED on this image (no Go toolchain).  API types of the AzureVmPool CRD: the field
// names, JSON tags and kubebuilder markers are the reference's contract (README.md:83-156) and
// are kept as they are so existing manifests (README.md:267-286) keep working.
package v1alpha1

import (
	metav1 "k8s.io/apimachinery/pkg/apis/meta/v1"
)

// ImageReferenceSpec names the VM image (README.md:113-118).
type ImageReferenceSpec struct {
	Publisher string `json:"publisher"`
	Offer     string `json:"offer"`
	SKU       string `json:"sku"`
	Version   string `json:"version"`
}

// AzureVmPoolSpec is the desired state (README.md:92-110).
type AzureVmPoolSpec struct {
	// +kubebuilder:validation:Minimum=0
	Replicas              int32              `json:"replicas"`
	ResourceGroupName     string             `json:"resourceGroupName"`
	Location              string             `json:"location"`
	VMSize                string             `json:"vmSize"`
	VnetName              string             `json:"vnetName"`
	SubnetName            string             `json:"subnetName"`
	ImageReference        ImageReferenceSpec `json:"imageReference"`
	AzureCredentialSecret string             `json:"azureCredentialSecret"`
}

// AzureVmPoolStatus is the observed state (README.md:121-128).
type AzureVmPoolStatus struct {
	ReadyReplicas int32              `json:"readyReplicas,omitempty"`
	VMs           []string           `json:"vms,omitempty"`
	Conditions    []metav1.Condition `json:"conditions,omitempty" patchStrategy:"merge" patchMergeKey:"type"`
}

//+kubebuilder:object:root=true
//+kubebuilder:subresource:status
//+kubebuilder:printcolumn:name="Desired",type="integer",JSONPath=".spec.replicas"
//+kubebuilder:printcolumn:name="Ready",type="integer",JSONPath=".status.readyReplicas"

// AzureVmPool is the Schema for the azurevmpools API (README.md:136-142).
type AzureVmPool struct {
	metav1.TypeMeta   `json:",inline"`
	metav1.ObjectMeta `json:"metadata,omitempty"`

	Spec   AzureVmPoolSpec   `json:"spec,omitempty"`
	Status AzureVmPoolStatus `json:"status,omitempty"`
}

//+kubebuilder:object:root=true

// AzureVmPoolList contains a list of AzureVmPool (README.md:147-151).
type AzureVmPoolList struct {
	metav1.TypeMeta `json:",inline"`
	metav1.ListMeta `json:"metadata,omitempty"`
	Items           []AzureVmPool `json:"items"`
}

func init() {
	SchemeBuilder.Register(&AzureVmPool{}, &AzureVmPoolList{})
}

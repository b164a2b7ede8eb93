// UNCOMPILED: see README.md in this directory.
module github.com/my-org/azure-vm-operator

go 1.22

require (
	k8s.io/apimachinery v0.29.0
	sigs.k8s.io/controller-runtime v0.17.0
)

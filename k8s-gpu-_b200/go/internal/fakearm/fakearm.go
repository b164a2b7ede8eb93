// UNCOMPILED on this image (no Go toolchain).  In-memory stand-in for the Azure SDK: resource
// groups of tagged VMs in creation order, idempotent create/delete, programmable failures
// (README.md:237-240).  Mirrors ../../host/arm_client.hpp, which is compiled and tested.
package fakearm

import "errors"

const (
	ManagedByKey   = "managed-by"
	ManagedByValue = "azurevmpool-operator" // README.md:238
	OwnerKey       = "owner"                // "<namespace>-<name>"
)

type VM struct {
	Name *string
	Tags map[string]string
}

type Kind int

const (
	Client Kind = iota
	List
	Create
	Delete
)

type Client struct {
	groups     map[string][]VM
	failAt     [4]int
	Calls      [4]int
	clientFail bool // getAzureVMClient fails while set (README.md:183-185: requeue after 30 s)
}

// FailClient arms or disarms the failure of step 2 (building the Azure client from the Secret).
func (c *Client) FailClient(on bool) { c.clientFail = on }

// ClientError is what getAzureVMClient returns for this fake: nil, or the armed failure.
func (c *Client) ClientError() error {
	if c.clientFail {
		return errors.New("fake ARM: cannot build the Azure VM client")
	}
	return nil
}

func New() *Client { return &Client{groups: map[string][]VM{}} }

// FailAt makes the k-th upcoming call of that kind fail (1-based); 0 disarms.
func (c *Client) FailAt(kind Kind, k int) { c.failAt[kind] = k }

func (c *Client) tick(kind Kind) bool {
	c.Calls[kind]++
	if c.failAt[kind] <= 0 {
		return false
	}
	c.failAt[kind]--
	return c.failAt[kind] == 0
}

func (c *Client) ListVMs(rg string) ([]VM, error) {
	if c.tick(List) {
		return nil, errors.New("ARM list failed (injected)")
	}
	return append([]VM(nil), c.groups[rg]...), nil
}

func (c *Client) CreateVM(rg string, vm VM) error {
	if c.tick(Create) {
		return errors.New("ARM create failed (injected)")
	}
	for _, v := range c.groups[rg] {
		if *v.Name == *vm.Name {
			return nil // create-if-absent, README.md:240
		}
	}
	c.groups[rg] = append(c.groups[rg], vm)
	return nil
}

// DeleteVM removes the VM with its NIC and OS disk (README.md:239); a missing VM is not an error.
func (c *Client) DeleteVM(rg, name string) error {
	if c.tick(Delete) {
		return errors.New("ARM delete failed (injected)")
	}
	g := c.groups[rg]
	for i := range g {
		if *g[i].Name == name {
			c.groups[rg] = append(g[:i:i], g[i+1:]...)
			break
		}
	}
	return nil
}

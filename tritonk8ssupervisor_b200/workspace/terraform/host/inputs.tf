# Inputs of the "host" module: one Kubernetes node.
# The generated terraform/rancher.tf sets hostname, networks, root_authorized_keys
# and package on every module block; image always falls back to the default below.
# (interface: reference terraform/host/vars.tf:1-23)

variable "hostname" {
  description = "name of the machine; also written to its tags"
}

variable "networks" {
  description = "UUIDs of the Triton networks the machine is attached to"
  type        = "list"
}

variable "package" {
  description = "Triton package (instance size): name or UUID"
  default     = "k4-highcpu-kvm-1.75G"
}

variable "image" {
  description = "image UUID; the default is ubuntu-certified-16.04"
  default     = "0867ef86-e69d-4aaa-ba3b-8d2aef0c204e"
}

variable "root_authorized_keys" {
  description = "key material used for root login and for the bootstrap connection"
  default     = "~/.ssh/id_rsa"
}

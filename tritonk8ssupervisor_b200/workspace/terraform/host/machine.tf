# one Kubernetes node: one KVM instance, bootstrapped just far enough for Ansible to take
# over (root reachable over ssh, a python interpreter present), its address
# appended to terraform/hosts.ip for the inventory generator.
# (behaviour: reference terraform/host/main.tf:1-36)

resource "triton_machine" "host" {
  name                 = "${var.hostname}"
  image                = "${var.image}"
  package              = "${var.package}"
  networks             = "${var.networks}"
  root_authorized_keys = "${var.root_authorized_keys}"

  tags = {
    name     = "${var.hostname}"
    hostname = "${var.hostname}"
  }

  connection {
    type        = "ssh"
    host        = "${self.primaryip}"
    user        = "ubuntu"
    private_key = "${var.root_authorized_keys}"
    agent       = false
  }

  # the image only lets "ubuntu" in; give root the same keys, then install the
  # interpreter Ansible's modules need on the target
  provisioner "remote-exec" {
    inline = [
      "sleep 30",
      "sudo install -m 600 -o root -g root /home/ubuntu/.ssh/authorized_keys /root/.ssh/authorized_keys",
      "sudo apt-get update",
      "sudo apt-get install -y python-minimal",
    ]
  }

  provisioner "local-exec" {
    command = "echo ${self.primaryip} >> hosts.ip"
  }
}

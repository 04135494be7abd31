"""HCL1 / HIL parser: the dialect of the generated rancher.tf and of the modules."""
import pytest

from tritonk8ssupervisor_b200 import hcl1
from tritonk8ssupervisor_b200.hcl1 import Body, HCall, HStr, HVar, ListVal


def test_generated_module_block():
    src = '''
module "kubenode1" {
    source = "host"
    hostname = "kubenode1"
    networks = ["a-1","b-2"]
    root_authorized_keys = "${file("~/.ssh/id_rsa")}"
    package = "14b6fade"
}
'''
    body = hcl1.parse(src)
    (mod,) = body.items
    assert mod.keys == ["module", "kubenode1"] and mod.is_block
    nets = mod.value.attr("networks").value
    assert isinstance(nets, ListVal) and [n.parts for n in nets.items] == [["a-1"], ["b-2"]]
    key = mod.value.attr("root_authorized_keys").value
    (call,) = key.parts
    assert isinstance(call, HCall) and call.func == "file"
    assert isinstance(call.args[0], HStr) and call.args[0].parts == ["~/.ssh/id_rsa"]


def test_comments_heredoc_numbers_bools_maps():
    src = '''
# hash comment
// slash comment
/* block
   comment */
variable "x" {
  default = 3
  flag = true
  ratio = 1.5
  hexv = 0x1F
  tags = { "name" = "v", plain = "w" }
  text = <<EOT
line one
  line two
EOT
  ind = <<-EOT
      a
        b
      EOT
}
'''
    v = hcl1.parse(src).items[0].value
    assert v.attr("default").value.value == 3
    assert v.attr("flag").value.value is True
    assert v.attr("ratio").value.value == 1.5
    assert v.attr("hexv").value.value == 31
    tags = v.attr("tags").value
    assert isinstance(tags, Body) and [i.name for i in tags.items] == ["name", "plain"]
    assert v.attr("text").value.value == "line one\n  line two\n"
    assert v.attr("ind").value.value == "a\n  b\n"


def test_repeated_attribute_kept_in_order():
    body = hcl1.parse('resource "t" "n" {\n tags = { a = "1" }\n x = "y"\n tags = { b = "2" }\n}')
    res = body.items[0].value
    assert [i.name for i in res.items] == ["tags", "x", "tags"]
    assert len(res.attrs("tags")) == 2


def test_hil_expressions():
    parts = hcl1.parse_template('pre ${var.a}-${triton_machine.m.primaryip} ${count.index + 1} $${lit}')
    assert parts[0] == "pre "
    assert parts[1] == HVar(["var", "a"])
    assert parts[3] == HVar(["triton_machine", "m", "primaryip"])
    assert isinstance(parts[5], hcl1.HBin) and parts[5].op == "+"
    assert parts[6] == " ${lit}"
    (cond,) = hcl1.parse_template('${var.n > 1 ? join(",", var.l) : element(var.l, 0)}')
    assert isinstance(cond, hcl1.HCond) and isinstance(cond.then, HCall) and cond.then.func == "join"
    (idx,) = hcl1.parse_template('${var.l[0]}')
    assert isinstance(idx, hcl1.HIndex)
    (splat,) = hcl1.parse_template('${triton_machine.m.*.primaryip}')
    assert splat == HVar(["triton_machine", "m", "*", "primaryip"])


def test_nested_quotes_and_interpolation_in_interpolation():
    (call,) = hcl1.parse_template('${format("%s-${var.x}", "a}b")}')
    assert call.func == "format"
    assert call.args[0].parts == ["%s-", HVar(["var", "x"])]
    assert call.args[1].parts == ["a}b"]


@pytest.mark.parametrize("src,needle", [
    ('module "a" {', "closing"),
    ('x = "unterminated', "not terminated"),
    ('x = "${var.a"', "not closed"),
    ('x = "a\nb"', "newline"),
    ("x = [1, 2", "expected ',' or ']'"),
    ("x = bareword", "must be quoted"),
    ('module "a" = "b"', "nested object expected"),
    ("= 3", "identifier"),
    ('x = "${}"', "empty interpolation"),
    ('x = "${var.a +}"', "unexpected"),
    ('x = "\\q"', "escape"),
    ("}", "unexpected '}'"),
    ("/* open", "comment"),
    ("x = <<EOT\nno end", "heredoc"),
])
def test_syntax_errors(src, needle):
    with pytest.raises(hcl1.HCLError) as ei:
        hcl1.parse(src, "t.tf")
    assert needle in str(ei.value)
    assert str(ei.value).startswith("t.tf:")


def test_error_position():
    with pytest.raises(hcl1.HCLError) as ei:
        hcl1.parse('a = "1"\nb = "2"\nc = ?', "f.tf")
    assert ei.value.line == 3

"""Offline drop-in and conformance harness for cheapRoc/tritonK8ssupervisor.

No GPU code lives here, by design: BASELINE.json flags the reference as a tier
mismatch with no GPU graft target.

  python -m tritonk8ssupervisor_b200 [-c]        the setup path (replacement for ./setup.sh)
  python -m tritonk8ssupervisor_b200.tools ...   validate / plan / apply-mock / syntax-check / inventory / lint / dry-run

The path itself (supervisor, bashsem, textops, tritonenv, runtime, cli) needs
only the standard library; the checkers additionally need PyYAML and Jinja2.
"""
__version__ = "0.1.0"

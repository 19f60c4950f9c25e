"""Loads dae-cpp_b200/binding.py (the package directory name is not a Python identifier)."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_spec = importlib.util.spec_from_file_location("daecpp_b200_binding", os.path.join(ROOT, "dae-cpp_b200", "binding.py"))
binding = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(binding)

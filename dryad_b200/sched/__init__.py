"""Scheduling layer: native graph manager bindings, stage statistics, event reporting."""

"""Headless observability tools (job report, diagnosis) — the information content of JobBrowser."""

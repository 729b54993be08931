# -*- coding: utf-8 -*-
"""Runtime helpers (the reference ships this directory without an ``__init__``: SURVEY Q11)."""

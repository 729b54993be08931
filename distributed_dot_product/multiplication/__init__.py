# -*- coding: utf-8 -*-
"""Distributed products and their autograd wrappers."""

"""Data-parallel machinery: process groups, flat buckets, codecs, all-reduce back-ends."""

"""Workload definitions ("model families" of a dataflow engine): TeraSort, GroupBy-aggregate,
WordCount, joins, k-means — each built on the public query API and the parallel substrate."""

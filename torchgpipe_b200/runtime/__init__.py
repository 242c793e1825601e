"""Pipeline runtime: schedule, persistent stage executors, streams, timeline."""

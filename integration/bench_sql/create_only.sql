.timer on
SET threads=16;
CREATE TABLE t AS SELECT i AS id, [((hash(i * 128 + x) % 100000) / 100000.0)::FLOAT for x in range(128)]::FLOAT[128] AS vec FROM range(1000000) r(i);
.print ##create_index
CREATE INDEX idx ON t USING HNSW (vec) WITH (metric = 'l2sq');

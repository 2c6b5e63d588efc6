-- BASELINE config 4: 5-way star join on TPC-H-shaped data with cascaded predicate transfer across the join graph.
-- Data: DuckDB's own dbgen (tpch extension, built into the shell); the scale factor is set by the runner
-- (SF100 needs ~100 GB of tables; the runner uses what the box holds — SF1 here, SF10 on the GPU box).
SELECT 'q_star5', n_name, count(*), sum(l_extendedprice * (1 - l_discount))::DECIMAL(18,2)
FROM lineitem JOIN orders ON l_orderkey = o_orderkey
              JOIN customer ON o_custkey = c_custkey
              JOIN nation ON c_nationkey = n_nationkey
              JOIN part ON l_partkey = p_partkey
WHERE o_orderdate >= DATE '1994-01-01' AND o_orderdate < DATE '1995-01-01'
  AND p_size < 10 AND n_regionkey = 2
GROUP BY n_name ORDER BY n_name;
SELECT 'q_star5_supp', count(*), sum(l_quantity)
FROM lineitem JOIN supplier ON l_suppkey = s_suppkey
              JOIN orders ON l_orderkey = o_orderkey
              JOIN customer ON o_custkey = c_custkey
              JOIN part ON l_partkey = p_partkey
WHERE s_acctbal > 9000 AND c_mktsegment = 'BUILDING' AND p_brand = 'Brand#13';
-- TPC-H Q3, Q10 (multi-join queries of the benchmark itself)
SELECT 'tpch_q3', count(*), sum(revenue)::DECIMAL(18,2) FROM (
  SELECT l_orderkey, sum(l_extendedprice * (1 - l_discount)) AS revenue, o_orderdate, o_shippriority
  FROM customer, orders, lineitem
  WHERE c_mktsegment = 'BUILDING' AND c_custkey = o_custkey AND l_orderkey = o_orderkey
    AND o_orderdate < DATE '1995-03-15' AND l_shipdate > DATE '1995-03-15'
  GROUP BY l_orderkey, o_orderdate, o_shippriority);
-- (TPC-H Q5 is left out: the UNMODIFIED reference extension aborts on it at plan time — "INTERNAL Error: Attempted to
--  access index 2 within vector of size 2" from its optimizer, with the cyclic customer/supplier nation conditions —
--  before any bloom filter exists; reproduced with baseline/_ref/bin/stock on the CPU box.)
SELECT 'tpch_q10', count(*), sum(revenue)::DECIMAL(18,2) FROM (
  SELECT c_custkey, sum(l_extendedprice * (1 - l_discount)) AS revenue
  FROM customer, orders, lineitem, nation
  WHERE c_custkey = o_custkey AND l_orderkey = o_orderkey AND o_orderdate >= DATE '1993-10-01'
    AND o_orderdate < DATE '1994-01-01' AND l_returnflag = 'R' AND c_nationkey = n_nationkey
  GROUP BY c_custkey);

"""The parameter-server role in full.  (Basics-Tutorial/Parameter-Server.ipynb)

A ps does nothing but host variables: start the server and join.  In the reference the
notebook kernel "stalls until it is shut down"; here join() returns once every worker
has reported that it is done (pass linger=True for the never-ending behaviour)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dtg_b200 as dtg

cluster_spec = dtg.ClusterSpec({"ps": ["localhost:2222"], "worker": ["localhost:2223"]})
ps = dtg.Server(cluster_spec, job_name="ps")
print("ps hosting variables at", ps.target)
ps.join()

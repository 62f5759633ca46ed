"""ps for the two-worker walkthrough.  (Basics-Tutorial/Multiple-Workers/Parameter-Server.ipynb)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import dtg_b200 as dtg

cluster_spec = dtg.ClusterSpec({"ps": ["localhost:2222"],
                                "worker": ["localhost:2223", "localhost:2224"]})
dtg.Server(cluster_spec, job_name="ps").join()

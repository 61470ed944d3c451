"""CPU oracle for the YOLOv1 head hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this package.  See yolo_oracle.c for the citations.
"""

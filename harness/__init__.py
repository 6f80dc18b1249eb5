"""Benchmark / integration harness around the correlation path (NOT part of the product).

SURVEY.md §2 rows 5-7 (RAFT glue, encoders, update block) are out of scope for this build and
stay stock PyTorch/cuDNN.  /root/reference does not exist on the GPU box, so the harness
re-states that caller with plain torch.nn modules — parameter names match the reference's
state_dict so its weights load unchanged — and takes the correlation block as a plug-in.
"""

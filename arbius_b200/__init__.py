"""arbius_b200 — a Blackwell-native Arbius mining inference framework.

Sub-packages:
  protocol  byte-level contracts (CIDv0, keccak/ABI, seed, hydration, templates, config)
  chain     EngineSim (EngineV1 state machine), JSON-RPC client/facade, wallet
  miner     durable job queue, event handlers, multi-GPU task scheduler, operator RPC
  worker    Cog-contract prediction server backed by the native models
  codecs    deterministic PNG / MP4 writers
  models    SD-1.5 (Anything V3), Kandinsky 2.1, UNet3D (ZeroScope / DAMO), RVM
  ops       ctypes bindings to the sm_100a kernels in csrc/
  parallel  NCCL groups, peer-memory heap, fused compute+collective kernels
"""
__version__ = "0.1.0"

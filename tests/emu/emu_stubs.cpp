// TEST INFRASTRUCTURE ONLY — entry points of csrc/api.cu that the CPU emulation does not cover (tensor-core search, peer
// copies, bridging tiles, UMAP rows): present so that the ctypes table of sph_b200/capi.py binds, every one returns
// SPHB_ERR_UNSUPPORTED (5).  Declared without the public header on purpose: the argument lists do not matter here.
#define EMU_STUB(name) extern "C" __attribute__((visibility("default"))) int name() { return 5; }
EMU_STUB(sphb_renorm_l2)
EMU_STUB(sphb_knn_rows_noalloc_dev)
EMU_STUB(sphb_knn_last_overflow_rows)
EMU_STUB(sphb_knn_graph_prow_multi)
EMU_STUB(sphb_component_centers)
EMU_STUB(sphb_component_centers_dev)
EMU_STUB(sphb_bridge_components)
EMU_STUB(sphb_bridge_components_dev)
EMU_STUB(sphb_prow_exponential_graph)
EMU_STUB(sphb_prow_exponential_graph_dev)
EMU_STUB(sphb_prow_exponential_inplace)
EMU_STUB(sphb_prow_exponential_inplace_dev)

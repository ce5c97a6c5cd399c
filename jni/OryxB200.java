package com.cloudera.oryx.app.als.gpu;

/**
 * JNI view of include/oryx_b200.h (liboryx_b200.so), one native method per C entry point a Java host needs.
 * Every method throws on a non-zero status (see oryx_b200_jni.c throw_for_status):
 * IllegalArgumentException, IllegalStateException, SingularMatrixSolverException, UnsupportedOperationException.
 *
 * <p>Shipped as source: this build image has no JVM and no jni.h, so neither file in jni/ is compiled
 * here beyond a syntax check of the C side against a stand-in header (tests/test_jni_source.py).
 * Call sites in the reference that change are listed in INTEGRATION.md section 2.
 */
public final class OryxB200 {
  static { System.loadLibrary("oryx_b200_jni"); }   // links against liboryx_b200.so

  private OryxB200() {}

  public static final int DOT = 0, COSINE = 1, X = 0, Y = 1;
  public static final int MAX_HOW_MANY = 65536;

  /** LocalitySensitiveHash sizing rule (LocalitySensitiveHash.java:45-74): {numHashes, maxBitsDiffering}. */
  public static native int[] lshPlan(double sampleRate, int numCores);

  public static native long modelCreate(int features, boolean implicit, int numHashes, int maxBitsDiffering,
                                        float[] hashVectorsRowMajor, int device, long capacityHint);
  public static native void modelDestroy(long model);

  /** ALSServingModel.setItemVector / setUserVector for a batch of rows; returns the LSH partition per row. */
  public static native int[] itemsUpsert(long model, int which, int[] rows, float[] vecsRowMajor);
  public static native void itemsRemove(long model, int which, int[] rows);
  /** null if the row has no vector. */
  public static native float[] itemsGet(long model, int which, int row, int features);
  public static native void itemsCompact(long model, int which);

  public static native int lshIndex(long model, float[] vec);
  public static native int[] lshCandidates(long model, float[] vec);

  /** out[0] = rows (int[]), out[1] = scores (double[]), best first (ALSServingModel.topN). */
  public static native Object[] topN(long model, int kind, float[] query, int howMany, int[] excludedRows);
  /** topN for many request threads: concurrent callers are answered together by one batch pass. */
  public static native Object[] topNCoalesced(long model, int kind, float[] query, int howMany, int[] excludedRows);
  /** out[0] = int[nq * howMany] rows, out[1] = double[nq * howMany] scores, out[2] = int[nq] counts. */
  public static native Object[] topNBatch(long model, int kind, float[] queriesRowMajor, int nq, int howMany,
                                          int[] exclOffsets, int[] exclRows);
  public static native double[] scoresAll(long model, int kind, float[] query, int nRows);
  /** Scores of the given rows; NaN where a row has no vector. */
  public static native double[] scoresRows(long model, int kind, float[] query, int[] rows);
  /** VectorMath.dot of (user row, item row) pairs; 0 where either side is unknown (Estimate.java:62-71). */
  public static native double[] estimatePairs(long model, int[] userRows, int[] itemRows);

  public static native void knownItemsAdd(long model, int[] userRows, long[] begin, int[] itemRows);
  public static native int[] knownItemsGet(long model, int userRow);
  public static native void knownItemsRetain(long model, byte[] keepUsers, byte[] keepItems);
  public static native int[] knownCounts(long model, int which, int nRows);
  /** null when the user has no vector (the REST layer answers 404). */
  public static native Object[] recommend(long model, int userRow, int howMany, boolean considerKnownItems);
  public static native Object[] recommendCoalesced(long model, int userRow, int howMany, boolean considerKnownItems);
  public static native Object[] recommendBatch(long model, int[] userRows, int howMany, boolean considerKnownItems);

  /** VectorMath.transposeTimesSelf summed over the store (getYTYSolver / getXTXSolver); null if empty. */
  public static native double[] gram(long model, int which, int features);
  /** Gram of a store sharded over the box's GPUs: this shard's Gram summed with the other ranks' by one kernel over
   *  peer memory; every rank calls it (from its own thread) and receives the same matrix. */
  public static native double[] gramSharded(long model, long exchange, int which, int features);
  /** LinearSystemSolver.getSolver; throws SingularMatrixSolverException(apparentRank, message). */
  public static native long solverCreate(double[] gramRowMajor, int features, int device);
  public static native void solverDestroy(long solver);
  public static native float[] solveFToF(long solver, float[] b);
  /** ALSUtils.computeUpdatedXu per interaction; result[i] == null where the reference returns null. */
  public static native float[][] foldInBatch(long solver, boolean implicit, float[][] xuOrNull, float[][] yi,
                                             double[] values);
  public static native float[][] foldInRows(long solver, long model, int features, boolean implicit, int xuWhich,
                                            int[] xuRows, int[] yiRows, double[] values);

  public static native long exchangeCreate(int device, int world, int rank, int maxHowMany);
  public static native void exchangeConnectLocal(long[] exchangesInRankOrder);
  public static native void exchangeDestroy(long exchange);

  /** {rowsX, rowsY, slotsX, slotsY, deviceBytes, queries, kernelLaunches}. */
  /** One JVM per GPU: export this model's X and Y stores, map the other ranks' (oryx_peer_stores_*). */
  public static native long peerStoresCreate(long model, int world, int rank);
  /** ORYX_PEER_STORES_HANDLE_BYTES (512) bytes for the other processes. */
  public static native byte[] peerStoresHandle(long peers);
  /** world x 512 bytes in rank order. */
  public static native void peerStoresConnect(long peers, byte[] handlesInRankOrder);
  public static native void peerStoresDestroy(long peers);
  /** computeUpdatedXu over rows that live in any rank's stores (global row ids); null rows = no result. */
  public static native float[][] foldInRowsPeers(long solver, long peers, int features, long xuRowStride, long yiRowStride,
                                                 boolean implicit, int xuWhich, long[] xuRows, long[] yiRows, double[] values);
  /** ALSSpeedModelManager.buildUpdates aggregation (sort by timestamp, sum-with-NaN / last-wins per (user, item))
   *  on the model's GPU; returns the number of groups written to the three output arrays. */
  public static native long aggregateInteractions(long model, int[] userRows, int[] itemRows, double[] strengths,
                                                  long[] timestamps, boolean implicit, int[] outUserRows, int[] outItemRows,
                                                  double[] outStrengths);
  public static native long[] stats(long model);
}

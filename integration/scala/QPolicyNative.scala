// Binds include/qpolicy.h (libqpolicy.so) with java.lang.foreign.  Uncompiled here: no JDK in the build image.
import java.lang.foreign._
import java.lang.foreign.ValueLayout._
import java.lang.invoke.MethodHandle

object QPolicyNative {
  private val linker = Linker.nativeLinker()
  private val lib    = SymbolLookup.libraryLookup(System.mapLibraryName("qpolicy"), Arena.global())
  private def fn(name: String, fd: FunctionDescriptor): MethodHandle =
    linker.downcallHandle(lib.find(name).orElseThrow(), fd)

  /** qp_config (include/qpolicy.h) */
  val cfgLayout: StructLayout = MemoryLayout.structLayout(
    JAVA_INT.withName("window"), JAVA_INT.withName("hidden"), JAVA_FLOAT.withName("epsilon"),
    JAVA_FLOAT.withName("gamma"), JAVA_FLOAT.withName("lr"), JAVA_FLOAT.withName("bias"),
    JAVA_FLOAT.withName("step_div"), JAVA_FLOAT.withName("adagrad_init_acc"),
    JAVA_INT.withName("bookkeeping"), JAVA_INT.withName("device"), JAVA_LONG.withName("seed"))

  val configDefault: MethodHandle = fn("qp_config_default", FunctionDescriptor.of(JAVA_INT, ADDRESS))
  val create: MethodHandle        = fn("qp_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS))
  val destroy: MethodHandle       = fn("qp_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS))
  val selectAction: MethodHandle  = fn("qp_select_action",
    FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG))
  val updateQ: MethodHandle       = fn("qp_update_q",
    FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, JAVA_LONG, JAVA_LONG))
  val enqueueSelect: MethodHandle = fn("qp_enqueue_select",
    FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_FLOAT, JAVA_LONG, ADDRESS))
  val enqueueUpdate: MethodHandle = fn("qp_enqueue_update",
    FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_FLOAT, ADDRESS, JAVA_LONG))
  val flush: MethodHandle         = fn("qp_flush", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS))
  val takeAction: MethodHandle    = fn("qp_take_action", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS))
  val train: MethodHandle         = fn("qp_train",
    FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_LONG, ADDRESS, JAVA_LONG, JAVA_LONG,
                          JAVA_DOUBLE, JAVA_INT, ADDRESS))
  val portfolioStats: MethodHandle = fn("qp_env_portfolio_stats", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS))
  val setSnapshot: MethodHandle   = fn("qp_set_snapshot", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG))
  val lastError: MethodHandle     = fn("qp_last_error", FunctionDescriptor.of(ADDRESS))

  /** QP_EINVAL is what the reference throws as IllegalArgumentException inside `receive`. */
  def check(status: Int): Unit = status match {
    case 0 => ()
    case 1 => throw new IllegalArgumentException(errorText())
    case s => throw new RuntimeException(s"qpolicy status $s: ${errorText()}")
  }
  private def errorText(): String =
    lastError.invoke().asInstanceOf[MemorySegment].reinterpret(4096).getString(0)

  def newPolicy(arena: Arena): MemorySegment = {
    val cfg = arena.allocate(cfgLayout)
    check(configDefault.invoke(cfg).asInstanceOf[Int])
    val out = arena.allocate(ADDRESS)
    check(create.invoke(cfg, out).asInstanceOf[Int])
    out.get(ADDRESS, 0)
  }
}

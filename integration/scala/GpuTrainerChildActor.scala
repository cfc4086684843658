// TrainerChildActor with train() (TrainerChildActor.scala:64-71) running the whole fold on the GPU
// through qp_train; FSM, messages and the size guard are inherited.  The reference's own tests
// override `train` the same way (TrainerRouterActorSpec.scala:144-153).
import QPolicyNative._
import SharePriceGetter.StockDataResponse
import akka.actor.ActorRef
import java.lang.foreign.{Arena, MemorySegment}
import java.lang.foreign.ValueLayout._
import scala.concurrent.Future

class GpuTrainerChildActor(policyHandle: MemorySegment, policyActor: ActorRef, myBudget: Double, noOfStocks: Int, envId: Long)
    extends TrainerChildActor(policyActor, myBudget, noOfStocks) {

  override def train(stockData: StockDataResponse): Future[Double] = Future {
    val prices = stockData.sharePrices.toIndexedSeq.map(_._2.toFloat)       // date-sorted TreeMap values, .toFloat (:66)
    val a = Arena.ofConfined()
    try {
      val seriesOfEnv = a.allocateFrom(JAVA_INT, 0)
      val out = a.allocate(JAVA_DOUBLE)
      // fewer than window + 1 prices -> QP_EINVAL -> IllegalArgumentException, as at :70
      check(train.invoke(policyHandle, a.allocateFrom(JAVA_FLOAT, prices.toArray: _*), prices.size.toLong, 1L,
                         seriesOfEnv, 1L, envId, myBudget, noOfStocks, out).asInstanceOf[Int])
      out.get(JAVA_DOUBLE, 0)                                                // budget + shares * shareValue (:68)
    } finally a.close()
  }
}

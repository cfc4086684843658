// Optional replacement for the shared policy actor when many TrainerChildActors ask at once
// (TrainerRouterActor.scala:44, :60-66): messages are queued natively and a whole fan-out is answered
// by one GPU launch per kind.  Replies are sent after the flush, so the children's `ask` pattern
// (TrainerChildActor.scala:93, :97) is unchanged.
import QDecisionPolicyActor._
import QPolicyNative._
import akka.actor.{Actor, ActorLogging, ActorRef, Timers}
import java.lang.foreign.{Arena, MemorySegment}
import java.lang.foreign.ValueLayout._
import scala.collection.mutable
import scala.concurrent.duration._

class BatchingPolicyActor(expectedChildren: Int, maxDelay: FiniteDuration = 200.micros)
    extends Actor with ActorLogging with Timers {
  private case object Flush
  private val arena   = Arena.ofShared()
  private val handle  = newPolicy(arena)
  private val waitingSelect = mutable.ArrayBuffer.empty[(Long, ActorRef)]
  private val waitingUpdate = mutable.ArrayBuffer.empty[ActorRef]
  override def postStop(): Unit = { destroy.invoke(handle); arena.close() }

  private def floats(a: Arena, xs: IndexedSeq[Float]) = a.allocateFrom(JAVA_FLOAT, xs.toArray: _*)
  private def queued = waitingSelect.size + waitingUpdate.size

  override def receive: Receive = {
    case SelectionAction(state, step) =>
      val a = Arena.ofConfined()
      try {
        val ticket = a.allocate(JAVA_LONG)
        check(enqueueSelect.invoke(handle, floats(a, state), state.size.toLong, step,
                                   sender().path.uid.toLong, ticket).asInstanceOf[Int])
        waitingSelect += ((ticket.get(JAVA_LONG, 0), sender()))
      } finally a.close()
      schedule()
    case UpdateQ(state, reward, nextState) =>
      val a = Arena.ofConfined()
      try check(enqueueUpdate.invoke(handle, floats(a, state), state.size.toLong, reward,
                                     floats(a, nextState), nextState.size.toLong).asInstanceOf[Int])
      finally a.close()
      waitingUpdate += sender()
      schedule()
    case Flush => flushNow()
  }

  private def schedule(): Unit =
    if (queued >= expectedChildren) flushNow() else timers.startSingleTimer(Flush, Flush, maxDelay)

  private def flushNow(): Unit = if (queued > 0) {
    check(flush.invoke(handle, MemorySegment.NULL, MemorySegment.NULL).asInstanceOf[Int])
    val a = Arena.ofConfined()
    try {
      val act = a.allocate(JAVA_INT)
      waitingSelect.foreach { case (ticket, ref) =>
        check(takeAction.invoke(handle, ticket, act).asInstanceOf[Int])
        ref ! actions(act.get(JAVA_INT, 0))
      }
    } finally a.close()
    waitingUpdate.foreach(_ ! Updated)
    waitingSelect.clear(); waitingUpdate.clear()
  }
}

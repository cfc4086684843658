// Same companion object and messages as the reference (QDecisionPolicyActor.scala:16-35); the
// TensorFlow graph and session (:38-50, :79-83) are replaced by one native policy handle.
import QDecisionPolicyActor._
import QPolicyNative._
import akka.actor.{Actor, ActorLogging, Props}
import java.lang.foreign.{Arena, MemorySegment}
import java.lang.foreign.ValueLayout._

object QDecisionPolicyActor {
  val actions = Seq(Buy, Sell, Hold)
  val inputDim = 201 + 2
  def props = Props[QDecisionPolicyActor]

  sealed trait Action
  case object Buy extends Action
  case object Sell extends Action
  case object Hold extends Action

  sealed trait DecisionPolicy
  case class SelectionAction(currentState: IndexedSeq[Float], step: Float) extends DecisionPolicy
  case class UpdateQ(state: IndexedSeq[Float], reward: Float, nextState: IndexedSeq[Float]) extends DecisionPolicy
  case object Updated extends DecisionPolicy
}

class QDecisionPolicyActor extends Actor with ActorLogging {
  private val arena  = Arena.ofShared()
  private val handle = newPolicy(arena)
  override def postStop(): Unit = { destroy.invoke(handle); arena.close() }

  private def floats(a: Arena, xs: IndexedSeq[Float]): MemorySegment = a.allocateFrom(JAVA_FLOAT, xs.toArray: _*)

  override def receive: Receive = {
    // the size guards of the reference (:56-57, :64-65) live in the library: QP_EINVAL -> IllegalArgumentException
    case SelectionAction(state, step) =>
      val a = Arena.ofConfined()
      try {
        val act = a.allocate(JAVA_INT)
        check(selectAction.invoke(handle, floats(a, state), state.size.toLong, a.allocateFrom(JAVA_FLOAT, step),
                                  MemorySegment.NULL, act, MemorySegment.NULL, 1L).asInstanceOf[Int])
        sender() ! actions(act.get(JAVA_INT, 0))
      } finally a.close()
    case UpdateQ(state, reward, nextState) =>
      val a = Arena.ofConfined()
      try {
        check(updateQ.invoke(handle, floats(a, state), state.size.toLong, a.allocateFrom(JAVA_FLOAT, reward),
                             floats(a, nextState), nextState.size.toLong, 1L).asInstanceOf[Int])
        sender() ! Updated
      } finally a.close()
  }
}

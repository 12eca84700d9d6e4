ssh test@127.0.0.1 "kill -9 8963"
ssh test@127.0.0.1 "kill -9 8964"
ssh test@127.0.0.1 "kill -9 8965"
ssh test@127.0.0.1 "kill -9 8966"
kill -9 8873

ssh test@127.0.0.1 "kill -9 9431"
ssh test@127.0.0.1 "kill -9 9432"
ssh test@127.0.0.1 "kill -9 9433"
kill -9 8873

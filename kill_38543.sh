ssh test@127.0.0.1 "kill -9 9675"
ssh test@127.0.0.1 "kill -9 9676"
kill -9 8873
